CMD="python bench.py --steps 1 --warmup 3 --e2e-steps 1 --no-cpu --parts=counts --min-seconds 0"
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:counts_scatter_kernel<\(int\)1>' -c 1 -f -o gpurun_out/r02v_cov_full $CMD > gpurun_out/r02v_cov_ncu.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:counts_scatter_kernel<\(int\)0>' -c 1 -f -o gpurun_out/r02v_dep_full $CMD > gpurun_out/r02v_dep_ncu.log 2>&1
ls -la gpurun_out/*.ncu-rep
