# Round-2 profiling pass (run under gpurun, ONE GPU; numbers printed by a run under ncu are never bench values):
#   1. launch list of the bench command (per-launch gpu__time_duration, serialised, cold cache -> shares only)
#   2. ncu --set full captures of the dominant kernels of one ELBO+grad evaluation
set -x
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-secondary --no-config5 --no-config4"
$B > gpurun_out/ncu_plain_run.json 2> gpurun_out/ncu_plain_run.err || exit 1      # must exit 0 without ncu first
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2_launches.csv $B > gpurun_out/ncu_launches.log 2>&1
N="ncu --set full --clock-control none --import-source on --kernel-name-base demangled"
# matched launches per evaluation: oz_gemm x10 (ZU, then per slab X(UC)^T, K Mfull^T, Sigma2, X^T R, then GC)
$N -k regex:oz_gemm_kernel -s 20 -c 5 -o gpurun_out/r2_oz_gemm -f $B > gpurun_out/ncu_ozgemm.log 2>&1
$N -k regex:oz_slice_cols_kernel -s 10 -c 3 -o gpurun_out/r2_oz_slice_cols -f $B > gpurun_out/ncu_slicecols.log 2>&1
$N -k regex:sep_quad_full_kernel -s 4 -c 1 -o gpurun_out/r2_sep_quad -f $B > gpurun_out/ncu_sepquad.log 2>&1
$N -k regex:acos_ -s 8 -c 2 -o gpurun_out/r2_acos -f $B > gpurun_out/ncu_acos.log 2>&1
ls -la gpurun_out/*.ncu-rep
