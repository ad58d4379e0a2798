# end-of-round evidence with the final build: parity rates on a larger sample, divergence curves, GAE sweep
mkdir -p gpurun_out/r2av
timeout 1500 python tools/parity_rates.py 1024 60 > gpurun_out/r2av/parity_rates_1024.txt 2>&1; cut -c1-200 gpurun_out/r2av/parity_rates_1024.txt
timeout 600 python tools/divergence_curves.py > gpurun_out/r2av/divergence_curves.txt 2>gpurun_out/r2av/div.err; head -n 4 gpurun_out/r2av/divergence_curves.txt | cut -c1-130
timeout 900 python tools/gae_sweep.py > gpurun_out/r2av/gae_sweep.txt 2>gpurun_out/r2av/sweep.err; tail -n 2 gpurun_out/r2av/gae_sweep.txt
