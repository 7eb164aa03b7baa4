set -x
nproc; nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
python -c "import emcee; print('emcee', emcee.__version__)" 2>&1 | tail -1
python -c "import h5py; print('h5py', h5py.__version__)" 2>&1 | tail -1
python -c "import silx; print('silx', silx.version)" 2>&1 | tail -1
python -c "import tables; print('tables', tables.__version__)" 2>&1 | tail -1
python tools/sanitize_target.py > gpurun_out/r02_sanitize_plain.log 2>&1; echo "plain rc $?"
tail -3 gpurun_out/r02_sanitize_plain.log
for wl in C1 C3A; do BI_B200_LIB=build/lib_timing.so python bench.py --no-cpu --workload $wl --steps 3 --warmup 3 > gpurun_out/r02_timing_$wl.json 2> gpurun_out/r02_timing_$wl.err; tail -2 gpurun_out/r02_timing_$wl.err; done
BI_B200_LIB=build/lib_timing.so python bench.py --no-cpu --workload C3A --walkers 148 --steps 3 --warmup 3 > gpurun_out/r02_timing_C3A_148.json 2> gpurun_out/r02_timing_C3A_148.err; tail -1 gpurun_out/r02_timing_C3A_148.err
timeout 600 compute-sanitizer --tool memcheck --log-file gpurun_out/r02_sanitizer_memcheck.txt python tools/sanitize_target.py > gpurun_out/r02_sanitize_memcheck_run.log 2>&1; echo "memcheck rc $?"
tail -5 gpurun_out/r02_sanitizer_memcheck.txt
