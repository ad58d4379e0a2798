mkdir -p gpurun_out/r2af
timeout 900 python tools/divergence_curves.py > gpurun_out/r2af/divergence_curves.txt 2> gpurun_out/r2af/divergence.err; echo "divergence rc=$?"; head -n 12 gpurun_out/r2af/divergence_curves.txt | cut -c1-140
timeout 1500 python tools/gae_sweep.py > gpurun_out/r2af/gae_sweep.txt 2> gpurun_out/r2af/gae_sweep.err; echo "sweep rc=$?"; tail -n 8 gpurun_out/r2af/gae_sweep.txt; tail -n 3 gpurun_out/r2af/gae_sweep.err
