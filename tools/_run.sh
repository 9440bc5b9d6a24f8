bash tools/gpu_round.sh r1j step_kernel
python tools/phase_trace.py > gpurun_out/phase_trace_r1j.json
python tools/overlap_trace.py > gpurun_out/overlap_trace_r1j.json
