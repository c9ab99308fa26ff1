# round 2 re-entry: GPU parity on the whole tree, smoke, bench (both arms), ncu launch list + full capture, modal lines
tag=${1:-r02g}; out=gpurun_out; mkdir -p $out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,memory.total --format=csv > $out/${tag}_gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q --durations=8 > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $out/${tag}_pytest_gpu.log
tail -15 $out/${tag}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 $out/${tag}_smoke.log
timeout 600 python bench.py --steps 20 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; rc=$?; echo "bench exit $rc"; tail -c 2500 $out/${tag}_bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err; echo "ref exit $?"; tail -c 800 $out/${tag}_bench_ref.json
if [ $rc -eq 0 ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_list.log 2>&1; echo "ncu list exit $?"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'face_flux|cell_update' \
    --launch-skip 12 --launch-count 3 -o $out/${tag}_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline \
    > $out/${tag}_ncu_full.log 2>&1; echo "ncu full exit $?"
fi
timeout 400 python bench.py --scheme modal --steps 10 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_modal.json 2> $out/${tag}_bench_modal.err; echo "modal exit $?"
timeout 400 python bench.py --scheme modal --limiter eigen --ic sod --steps 10 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_modal_eigen.json 2> $out/${tag}_bench_modal_eigen.err; echo "modal eigen exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'modal_' \
    --launch-skip 10 --launch-count 8 -o $out/${tag}_modal_full -f python bench.py --scheme modal --limiter eigen --ic sod --steps 2 --warmup 3 --no-cpu-baseline \
    > $out/${tag}_ncu_modal_full.log 2>&1; echo "ncu modal full exit $?"
cut -c1-400 $out/${tag}_bench_modal.json $out/${tag}_bench_modal_eigen.json
