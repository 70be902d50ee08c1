# One --set full capture each of the two counting kernels (100 M-read launch) : bash profiles/prof_counts.sh TAG
TAG=${1:-r02w}
CMD="python bench.py --steps 1 --warmup 3 --e2e-steps 1 --no-cpu --parts=counts --min-seconds 0"
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:coverage_scatter_kernel' -c 1 -f -o gpurun_out/${TAG}_cov_full $CMD > gpurun_out/${TAG}_cov_ncu.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:depth_window_kernel' -c 1 -f -o gpurun_out/${TAG}_dep_full $CMD > gpurun_out/${TAG}_dep_ncu.log 2>&1
ls -la gpurun_out/*.ncu-rep
