# usage (on the GPU box, via gpurun): bash scripts/gpu_round.sh <tag>
# GPU parity tests, smoke, bench (own arm + reference arm), then -- only after the plain run exited 0 --
# the ncu launch list and one full capture of the two hot kernels.  Everything lands in gpurun_out/<tag>_*.
tag=${1:-r01}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,memory.total --format=csv > $out/${tag}_gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $out/${tag}_pytest_gpu.log
tail -5 $out/${tag}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 $out/${tag}_smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; rc=$?; echo "bench exit $rc"; tail -c 3000 $out/${tag}_bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err; echo "ref exit $?"; tail -c 1500 $out/${tag}_bench_ref.json
if [ $rc -eq 0 ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_list.log 2>&1; echo "ncu list exit $?"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'face_flux|cell_update' \
    --launch-skip 12 --launch-count 4 -o $out/${tag}_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline \
    > $out/${tag}_ncu_full.log 2>&1; echo "ncu full exit $?"
fi
# extra workloads (bench lines only, no ncu): modal DG with and without the Eigen-WENO limiter, DGSEM, and the
# 128^3 p=3 single-GPU slice of configs[4]
timeout 400 python bench.py --scheme modal --steps 10 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_modal.json 2> $out/${tag}_bench_modal.err; echo "modal exit $?"
timeout 400 python bench.py --scheme modal --limiter eigen --ic sod --steps 10 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_modal_eigen.json 2> $out/${tag}_bench_modal_eigen.err; echo "modal eigen exit $?"
timeout 400 python bench.py --scheme dgsem --steps 10 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_dgsem.json 2> $out/${tag}_bench_dgsem.err; echo "dgsem exit $?"
timeout 600 python bench.py --n 128 --degree 3 --steps 3 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_128_p3.json 2> $out/${tag}_bench_128_p3.err; echo "128^3 p3 exit $?"
tail -c 600 $out/${tag}_bench_modal.json $out/${tag}_bench_modal_eigen.json $out/${tag}_bench_dgsem.json $out/${tag}_bench_128_p3.json
