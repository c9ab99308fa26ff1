# round 2, first call: GPU parity on the new tests, bench baseline, FP64 peak, full ncu capture of both hot kernels
tag=${1:-r02a}; out=gpurun_out; mkdir -p $out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,memory.total --format=csv > $out/${tag}_gpu.txt 2>&1
nproc > $out/${tag}_nproc.txt
timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $out/${tag}_pytest_gpu.log
tail -15 $out/${tag}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 $out/${tag}_smoke.log
timeout 120 minicfd_b200/lib/tools/fp64_peak > $out/${tag}_fp64_peak.json 2>&1; cat $out/${tag}_fp64_peak.json
timeout 600 python bench.py --steps 20 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; rc=$?; echo "bench exit $rc"; tail -c 2500 $out/${tag}_bench.json
if [ $rc -eq 0 ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'face_flux_kernel|cell_update' \
    --launch-skip 12 --launch-count 3 -o $out/${tag}_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline \
    > $out/${tag}_ncu_full.log 2>&1; echo "ncu full exit $?"
fi
