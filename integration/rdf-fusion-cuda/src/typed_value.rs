//! ENC_TV on the device: the typed-value table next to the dictionary, value FILTERs evaluated on object ids, and the decode
//! into the TypedValue union.  NOT COMPILED in the repository this file lives in (no Rust toolchain there); see README.md.
//!
//! Mirrors lib/storage/src/memory/object_id_mapping.rs:376-402 (decode_array_to_typed_value) and the comparison operators of
//! lib/functions/src/scalar/comparison/*.rs.  The library never parses a lexical form: the records are filled from the
//! `EncodedTypedValue` the reference already keeps per interned term (lib/storage/src/memory/encoding/typed_value.rs:9-85).
use crate::{check, ffi, Gpu, GpuTable};
use datafusion::error::Result as DFResult;
use rdf_fusion_model::{Numeric, TypedValueRef};
use std::ffi::CString;
use std::ptr;
use std::sync::Arc;

fn i128_words(v: i128) -> (u64, u64) {
    (v as u64, ((v as u128) >> 64) as u64)
}

/// One `rfgpu_tv_value` (32 bytes) from a typed value; `None` (unbound, or a literal outside its datatype's lexical space:
/// EncodedTypedValue::Invalid) is the all-zero record = RFGPU_TV_NULL.
pub fn tv_record(value: Option<TypedValueRef<'_>>) -> ffi::rfgpu_tv_value {
    let mut r = ffi::rfgpu_tv_value { lo: 0, hi: 0, aux: 0, field: ffi::RFGPU_TV_NULL as u8, flags: 0, reserved: [0; 6] };
    let Some(value) = value else { return r };
    let mut dec = |r: &mut ffi::rfgpu_tv_value, d: rdf_fusion_model::Decimal| {
        (r.lo, r.hi) = i128_words(i128::from_be_bytes(d.to_be_bytes()));
    };
    match value {
        TypedValueRef::NamedNode(_) => r.field = ffi::RFGPU_TV_NAMED_NODE as u8,
        TypedValueRef::BlankNode(_) => r.field = ffi::RFGPU_TV_BLANK_NODE as u8,
        // the strings stay in the rfgpu_dict; whether a string carries a language tag is the dictionary's `language` index
        TypedValueRef::SimpleLiteral(_) | TypedValueRef::LanguageStringLiteral(_) => r.field = ffi::RFGPU_TV_STRING as u8,
        TypedValueRef::OtherLiteral(_) => r.field = ffi::RFGPU_TV_OTHER_LITERAL as u8,
        TypedValueRef::BooleanLiteral(b) => {
            r.field = ffi::RFGPU_TV_BOOLEAN as u8;
            r.lo = b.as_bool() as u64;
        }
        TypedValueRef::NumericLiteral(n) => match n {
            Numeric::Int(v) => (r.field, r.lo) = (ffi::RFGPU_TV_INT as u8, i32::from(v) as i64 as u64),
            Numeric::Integer(v) => (r.field, r.lo) = (ffi::RFGPU_TV_INTEGER as u8, i64::from(v) as u64),
            Numeric::Float(v) => (r.field, r.lo) = (ffi::RFGPU_TV_FLOAT as u8, f32::from(v).to_bits() as u64),
            Numeric::Double(v) => (r.field, r.lo) = (ffi::RFGPU_TV_DOUBLE as u8, f64::from(v).to_bits()),
            Numeric::Decimal(d) => {
                r.field = ffi::RFGPU_TV_DECIMAL as u8;
                dec(&mut r, d);
                // Double::from(Decimal) (lib/model/src/xsd/decimal.rs:445-462), precomputed: the device compares a decimal with a
                // float / double through it and never divides an i128
                r.aux = f64::from(rdf_fusion_model::Double::from(d)).to_bits() as i64;
            }
        },
        TypedValueRef::DateTimeLiteral(t) => timestamp(&mut r, ffi::RFGPU_TV_DATE_TIME, t.timestamp()),
        TypedValueRef::TimeLiteral(t) => timestamp(&mut r, ffi::RFGPU_TV_TIME, t.timestamp()),
        TypedValueRef::DateLiteral(t) => timestamp(&mut r, ffi::RFGPU_TV_DATE, t.timestamp()),
        TypedValueRef::DurationLiteral(d) => {
            r.field = ffi::RFGPU_TV_DURATION as u8;
            r.flags = (ffi::RFGPU_TV_HAS_AUX | ffi::RFGPU_TV_HAS_VALUE) as u8;
            r.aux = d.year_month().as_i64();
            dec(&mut r, d.day_time().as_seconds());
        }
        TypedValueRef::YearMonthDurationLiteral(d) => {
            r.field = ffi::RFGPU_TV_DURATION as u8;
            r.flags = ffi::RFGPU_TV_HAS_AUX as u8;
            r.aux = d.as_i64();
        }
        TypedValueRef::DayTimeDurationLiteral(d) => {
            r.field = ffi::RFGPU_TV_DURATION as u8;
            r.flags = ffi::RFGPU_TV_HAS_VALUE as u8;
            dec(&mut r, d.as_seconds());
        }
    }
    r
}

fn timestamp(r: &mut ffi::rfgpu_tv_value, field: i32, t: rdf_fusion_model::Timestamp) {
    r.field = field as u8;
    (r.lo, r.hi) = i128_words(i128::from_be_bytes(t.value().to_be_bytes()));
    if let Some(offset) = t.offset() {
        r.aux = offset.in_minutes() as i64;
        r.flags = ffi::RFGPU_TV_HAS_AUX as u8;
    }
}

/// The device twin of the per-id EncodedTypedValue table; refreshed (rebuilt) after loads, like the dictionary snapshot.
pub struct GpuTypedValues {
    gpu: Arc<Gpu>,
    raw: *mut ffi::rfgpu_tv_table,
}
unsafe impl Send for GpuTypedValues {}
unsafe impl Sync for GpuTypedValues {}
impl Drop for GpuTypedValues {
    fn drop(&mut self) {
        unsafe { ffi::rfgpu_tv_table_release(self.raw) }
    }
}

/// `col <op> constant` or `col <op> col`, the leaves of a value FILTER; operands index the input table's columns.
pub enum TvOperand {
    Column(i32),
    Constant(usize),
}
pub struct TvConstant<'a> {
    pub value: ffi::rfgpu_tv_value,
    pub string: &'a str,           // value of a named node / blank node / string / other literal ("" otherwise)
    pub lang_or_datatype: &'a str, // language tag (set RFGPU_TV_HAS_AUX in value.flags) or datatype IRI
}

impl GpuTypedValues {
    /// `records[id]` for id in 0..=n_ids (entry 0 unused), `dict` the rfgpu_dict over the same ids.
    pub fn new(gpu: Arc<Gpu>, dict: *mut ffi::rfgpu_dict, records: &[ffi::rfgpu_tv_value]) -> DFResult<Self> {
        let mut raw = ptr::null_mut();
        check(gpu.raw(), unsafe {
            ffi::rfgpu_tv_table_create(gpu.raw(), dict, records.as_ptr(), records.len() as u32 - 1, &mut raw)
        })?;
        Ok(Self { gpu, raw })
    }

    /// EBV(program) per row as a nullable id-level column `name` (1 true, 2 false, NULL error): FilterExec over
    /// `EBV(GT(ENC_TV(col), c))` becomes this call followed by `GpuIdFilterExec(name = 1)`.  RFGPU_E_UNSUPPORTED (duration
    /// order) surfaces as `DataFusionError::NotImplemented`: the optimizer rule then keeps the CPU FilterExec.
    pub fn eval(&self, input: &GpuTable, prog: &[(i32, Option<TvOperand>, Option<TvOperand>)], consts: &[TvConstant<'_>],
                name: &str) -> DFResult<GpuTable> {
        let operand = |o: &Option<TvOperand>| match o {
            Some(TvOperand::Column(c)) => *c,
            Some(TvOperand::Constant(k)) => -(*k as i32 + 1),
            None => 0,
        };
        let instrs: Vec<ffi::rfgpu_tv_instr> =
            prog.iter().map(|(op, l, r)| ffi::rfgpu_tv_instr { op: *op, lhs: operand(l), rhs: operand(r) }).collect();
        let cs: Vec<ffi::rfgpu_tv_const> = consts
            .iter()
            .map(|c| ffi::rfgpu_tv_const {
                v: c.value,
                str: c.string.as_ptr(),
                str_len: c.string.len() as u32,
                str2: c.lang_or_datatype.as_ptr(),
                str2_len: c.lang_or_datatype.len() as u32,
            })
            .collect();
        let name = CString::new(name).unwrap();
        let mut raw = ptr::null_mut();
        check(self.gpu.raw(), unsafe {
            ffi::rfgpu_tv_eval(self.gpu.raw(), self.raw, input.raw, instrs.as_ptr(), instrs.len() as i32, cs.as_ptr(),
                               cs.len() as i32, name.as_ptr(), &mut raw)
        })?;
        Ok(GpuTable { gpu: Arc::clone(&self.gpu), raw })
    }

    /// ENC_TV(col): the exported buffers are the buffers of the dense UnionArray of TypedValueEncoding
    /// (lib/encoding/src/typed_value/encoding.rs:16-137); assemble with `UnionArray::try_new(TypedValueEncoding::fields(),
    /// type_ids, Some(offsets), children)` as TypedValueArrayBuilder::finish does (builder.rs).  Left as a sketch: sizes ->
    /// allocate `Buffer`s -> rfgpu_typed_values_export / rfgpu_plain_terms_export for the four string children.
    pub fn decode_sizes(&self, input: &GpuTable, col: i32, row_off: u64, n_rows: u64) -> DFResult<(u64, [u64; 15])> {
        let mut raw = ptr::null_mut();
        check(self.gpu.raw(), unsafe { ffi::rfgpu_tv_decode(self.gpu.raw(), self.raw, input.raw, col, row_off, n_rows, &mut raw) })?;
        let (mut rows, mut counts) = (0u64, [0u64; 15]);
        let rc = unsafe { ffi::rfgpu_typed_values_sizes(raw, &mut rows, counts.as_mut_ptr()) };
        unsafe { ffi::rfgpu_typed_values_release(raw) };
        check(self.gpu.raw(), rc)?;
        Ok((rows, counts))
    }
}
