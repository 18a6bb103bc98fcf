set -x
for v in main m10 nopf; do
  if [ $v = main ]; then unset RFGPU_LIB; else export RFGPU_LIB=$PWD/.wip/librfgpu_$v.so; fi
  timeout 600 python bench.py --steps 10 --no-e2e --no-cpu-baseline > gpurun_out/r2z_bench_$v.json 2> gpurun_out/r2z_bench_$v.err
done
