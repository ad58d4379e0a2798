# round 2, run D: full GPU test pass + timing of the product build
mkdir -p gpurun_out/r2d
python -m pytest tests -m gpu -q > gpurun_out/r2d/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2d/pytest_gpu.log
python tools/phase_profile.py 160 60 > gpurun_out/r2d/time_base.txt 2>&1; tail -1 gpurun_out/r2d/time_base.txt
python tools/phase_profile.py 160 30 3 16384 > gpurun_out/r2d/time_3v3.txt 2>&1; tail -1 gpurun_out/r2d/time_3v3.txt
python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/r2d/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r2d/smoke.log
