timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r01q_pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/r01q_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 400 python bench.py --steps 10 --warmup 3 > gpurun_out/r01q_bench.json 2> gpurun_out/r01q_bench.err; echo "bench exit $?"; tail -c 1500 gpurun_out/r01q_bench.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01q_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r01q_ncu_list.log 2>&1; echo "ncu list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'face_flux_kernel|cell_update' --launch-skip 12 --launch-count 4 -o gpurun_out/r01q_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r01q_ncu_full.log 2>&1; echo "ncu full exit $?"
