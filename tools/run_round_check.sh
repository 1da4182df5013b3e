python -m pytest tests -m gpu -x -q > gpurun_out/r3e_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r3e_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r3e_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r3e_smoke.log
python bench.py > gpurun_out/r3e_bench.json 2> gpurun_out/r3e_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r3e_reference_arm.json 2>/dev/null; echo "ref rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r3e_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-latency --no-also > gpurun_out/r3e_ncu_launches.log 2>&1; echo "launches rc=$?"
PROF_B=32768 PROF_S=1 bash tools/ncu_capture.sh r3e "k_backward|k_forward|k_node_diff" 0 40 | tail -12
