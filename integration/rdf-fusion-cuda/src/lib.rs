//! Safe handles over the C ABI of librfgpu (include/rfgpu.h).  NOT COMPILED in the repository this file lives in (no Rust
//! toolchain there); see README.md.
pub mod data_source;
pub mod ffi;
pub mod join_exec;
pub mod optimizer;
pub mod planner;
pub mod typed_value;

use arrow::array::{Array, RecordBatch, StructArray};
use arrow::datatypes::SchemaRef;
use arrow::ffi::{from_ffi, to_ffi, FFI_ArrowArray, FFI_ArrowSchema};
use datafusion::common::{exec_err, not_impl_err, plan_err};
use datafusion::error::Result as DFResult;
use std::ffi::{CStr, CString};
use std::ptr;
use std::sync::Arc;

/// RFGPU_E_* -> DataFusionError, the `DFResult` convention of pattern_data_source.rs:47-55.  RFGPU_E_UNSUPPORTED tells the
/// optimizer rule to keep the CPU node.
pub(crate) fn check(ctx: *const ffi::rfgpu_ctx, rc: i32) -> DFResult<()> {
    if rc == ffi::RFGPU_OK {
        return Ok(());
    }
    let msg = unsafe { CStr::from_ptr(ffi::rfgpu_last_error(ctx)) }.to_string_lossy().into_owned();
    match rc {
        ffi::RFGPU_E_INVALID => plan_err!("{msg}"),
        ffi::RFGPU_E_UNSUPPORTED => not_impl_err!("{msg}"),
        _ => exec_err!("{msg}"),
    }
}

/// One CUDA device + stream.  Calls on one ctx must be serialised (rfgpu.h conventions): one `Gpu` per DataFusion
/// partition, used from `spawn_blocking` so that a tokio worker never waits on the device.
pub struct Gpu {
    ctx: *mut ffi::rfgpu_ctx,
}
unsafe impl Send for Gpu {}
unsafe impl Sync for Gpu {}

impl Gpu {
    pub fn new(device: i32) -> DFResult<Arc<Self>> {
        let mut ctx = ptr::null_mut();
        check(ptr::null(), unsafe { ffi::rfgpu_ctx_create(device, &mut ctx) })?; // RFGPU_E_NOGPU: there is no CPU fallback
        Ok(Arc::new(Self { ctx }))
    }
    pub(crate) fn raw(&self) -> *mut ffi::rfgpu_ctx {
        self.ctx
    }
}
impl Drop for Gpu {
    fn drop(&mut self) {
        unsafe { ffi::rfgpu_ctx_destroy(self.ctx) }
    }
}

/// IndexPermutations<MemQuadIndex> on the device; one retained snapshot per plan (snapshot.rs:35-37).
pub struct GpuStore {
    pub(crate) gpu: Arc<Gpu>,
    pub(crate) raw: *mut ffi::rfgpu_store,
}
unsafe impl Send for GpuStore {}
unsafe impl Sync for GpuStore {}
impl Clone for GpuStore {
    fn clone(&self) -> Self {
        Self { gpu: Arc::clone(&self.gpu), raw: unsafe { ffi::rfgpu_store_retain(self.raw) } }
    }
}
impl Drop for GpuStore {
    fn drop(&mut self) {
        unsafe { ffi::rfgpu_store_release(self.raw) }
    }
}
impl GpuStore {
    /// Bulk load: id quads (g, s, p, o columns, any order, duplicates allowed) -> the reference's three permutations plus
    /// GPSO, sorted and de-duplicated on the device (replaces MemQuadIndex::insert: quad_index.rs:132-156).
    pub fn build(gpu: Arc<Gpu>, quads: [&[u32]; 4]) -> DFResult<Self> {
        let mut raw = ptr::null_mut();
        check(gpu.raw(), unsafe { ffi::rfgpu_store_create(gpu.raw(), &mut raw) })?;
        let cols = [quads[0].as_ptr(), quads[1].as_ptr(), quads[2].as_ptr(), quads[3].as_ptr()];
        let perms = [
            ffi::rfgpu_perm(0, 1, 2, 3), // GSPO
            ffi::rfgpu_perm(0, 2, 3, 1), // GPOS
            ffi::rfgpu_perm(0, 3, 1, 2), // GOSP
            ffi::rfgpu_perm(0, 2, 1, 3), // GPSO: lets `?s <p> ?o` come out sorted on ?s (merge joins)
        ];
        let rc = unsafe {
            ffi::rfgpu_store_build(gpu.raw(), raw, cols.as_ptr(), quads[0].len() as u64, 0, perms.as_ptr(), perms.len() as i32)
        };
        let store = Self { gpu, raw };
        check(store.gpu.raw(), rc)?;
        Ok(store)
    }
}

/// A device-resident set of solution mappings.  Adjacent GPU nodes pass these instead of RecordBatches, so a
/// scan -> join -> join -> filter chain never leaves HBM; only the top of a GPU subtree exports Arrow.
pub struct GpuTable {
    pub(crate) gpu: Arc<Gpu>,
    pub(crate) raw: *mut ffi::rfgpu_table,
}
unsafe impl Send for GpuTable {}
unsafe impl Sync for GpuTable {}
impl Drop for GpuTable {
    fn drop(&mut self) {
        unsafe { ffi::rfgpu_table_release(self.raw) }
    }
}
impl GpuTable {
    pub fn num_rows(&self) -> u64 {
        unsafe { ffi::rfgpu_table_num_rows(self.raw) }
    }
    /// RecordBatch of UInt32 columns -> device (Arrow C Data Interface; validity bitmaps are folded into id 0 = NULL).
    pub fn import(gpu: Arc<Gpu>, batch: &RecordBatch) -> DFResult<Self> {
        let data = StructArray::from(batch.clone()).into_data();
        let (array, schema) = to_ffi(&data)?;
        let mut raw = ptr::null_mut();
        check(gpu.raw(), unsafe { ffi::rfgpu_table_import_arrow(gpu.raw(), &array, &schema, &mut raw) })?;
        Ok(Self { gpu, raw })
    }
    /// Rows [row_off, row_off + n_rows) as a RecordBatch with the node's schema (field names = SPARQL variables; graph and
    /// OPTIONAL-side columns nullable: quad_storage_encoding.rs:79-93, join/logical.rs:251-279).
    pub fn export(&self, schema: &SchemaRef, row_off: u64, n_rows: u64) -> DFResult<RecordBatch> {
        let (mut array, mut ffi_schema) = (FFI_ArrowArray::empty(), FFI_ArrowSchema::empty());
        check(self.gpu.raw(), unsafe {
            ffi::rfgpu_table_export_arrow(self.gpu.raw(), self.raw, row_off, n_rows, &mut array, &mut ffi_schema)
        })?;
        let data = unsafe { from_ffi(array, &ffi_schema) }?;
        let batch = RecordBatch::from(StructArray::from(data));
        Ok(batch.with_schema(Arc::clone(schema))?)
    }
}

/// PlannedPatternScan (scan.rs:414-513) on the device.
pub struct GpuScanPlan {
    pub(crate) store: GpuStore,
    pub(crate) raw: *mut ffi::rfgpu_scan_plan,
    pub(crate) is_empty: bool,
}
unsafe impl Send for GpuScanPlan {}
unsafe impl Sync for GpuScanPlan {}
impl Drop for GpuScanPlan {
    fn drop(&mut self) {
        unsafe { ffi::rfgpu_scan_plan_destroy(self.raw) }
    }
}
impl GpuScanPlan {
    /// "[GPOS] subject=?s, ..., additional_filters=[object in (2..9)]" (pattern_data_source.rs:89-104 fmt_as)
    pub fn describe(&self) -> String {
        let n = unsafe { ffi::rfgpu_scan_plan_describe(self.raw, ptr::null_mut(), 0) };
        let mut buf = vec![0u8; n];
        unsafe { ffi::rfgpu_scan_plan_describe(self.raw, buf.as_mut_ptr() as *mut _, n) };
        CStr::from_bytes_until_nul(&buf).map(|s| s.to_string_lossy().into_owned()).unwrap_or_default()
    }
    pub fn apply_filter(&mut self, column: &str, op: i32, value: u32) -> DFResult<()> {
        let c = CString::new(column).unwrap();
        check(self.store.gpu.raw(), unsafe {
            ffi::rfgpu_scan_plan_apply_filter(self.store.gpu.raw(), self.raw, c.as_ptr(), op, value)
        })
    }
    /// Runs the scan; `allow_view`: unfiltered ranges come back as zero-copy views of the index (scan.rs:146-170).
    pub fn execute(&self, allow_view: bool) -> DFResult<GpuTable> {
        let mut raw = ptr::null_mut();
        let flags = if allow_view { 1 } else { 0 };
        check(self.store.gpu.raw(), unsafe { ffi::rfgpu_scan_execute(self.store.gpu.raw(), self.raw, flags, &mut raw) })?;
        Ok(GpuTable { gpu: Arc::clone(&self.store.gpu), raw })
    }
}
