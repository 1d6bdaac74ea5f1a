# ncu --set full captures of the dominant kernels of one ELBO+grad evaluation (run under gpurun, one GPU)
set -x
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
N="ncu --set full --clock-control none --import-source on --kernel-name-base demangled"
$N -k "regex:gemm_f64_kernel<\(int\)64, \(int\)128, \(int\)0, \(int\)0, \(int\)0" -s 6 -c 1 -o gpurun_out/xuc_v4 -f $B > gpurun_out/ncu_xuc.log 2>&1
$N -k "regex:gemm_f64_kernel<\(int\)64, \(int\)128, \(int\)1, \(int\)1, \(int\)0, \(bool\)0" -s 6 -c 1 -o gpurun_out/xtr_v4 -f $B > gpurun_out/ncu_xtr.log 2>&1
ls -la gpurun_out/*.ncu-rep
