# final tree of a round: GPU parity, smoke, bench (both arms), ncu launch list + full captures, reference arm of the other configs
tag=${1:-r02u}; out=gpurun_out; mkdir -p $out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,memory.total --format=csv > $out/${tag}_gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q --durations=5 > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $out/${tag}_pytest_gpu.log; tail -4 $out/${tag}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke exit $?"; tail -1 $out/${tag}_smoke.log
timeout 600 python bench.py --steps 20 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; rc=$?; echo "bench exit $rc"; cut -c1-400 $out/${tag}_bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err; echo "ref exit $?"
if [ $rc -eq 0 ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_list.log 2>&1; echo "ncu list exit $?"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'face_flux|cell_update' \
    --launch-skip 8 --launch-count 2 -o $out/${tag}_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline \
    > $out/${tag}_ncu_full.log 2>&1; echo "ncu full exit $?"
fi
timeout 400 python bench.py --scheme modal --steps 10 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_modal.json 2> $out/${tag}_bench_modal.err; echo "modal exit $?"
timeout 400 python bench.py --scheme modal --limiter eigen --ic sod --steps 10 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_modal_eigen.json 2> $out/${tag}_bench_modal_eigen.err; echo "modal eigen exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'modal_' \
    --launch-skip 8 --launch-count 7 -o $out/${tag}_modal_full -f python bench.py --scheme modal --limiter eigen --ic sod --steps 2 --warmup 3 --no-cpu-baseline \
    > $out/${tag}_ncu_modal_full.log 2>&1; echo "ncu modal full exit $?"
timeout 600 python bench.py --impl reference --workload double_mach --scheme modal --limiter eigen --cells 120 --dt 5e-5 --steps 2 --warmup 1 > $out/${tag}_bench_ref_dmach.json 2> $out/${tag}_bench_ref_dmach.err; echo "ref dmach exit $?"
timeout 600 python bench.py --impl reference --workload forward_step --scheme modal --limiter eigen --cells 80 --dt 2e-5 --steps 2 --warmup 1 > $out/${tag}_bench_ref_fstep.json 2> $out/${tag}_bench_ref_fstep.err; echo "ref fstep exit $?"
timeout 600 python bench.py --impl reference --cells 256 --degree 3 --steps 2 --warmup 1 > $out/${tag}_bench_ref_256p3.json 2> $out/${tag}_bench_ref_256p3.err; echo "ref 256p3 exit $?"
for f in ref ref_dmach ref_fstep ref_256p3 modal modal_eigen; do python - $out/${tag}_bench_$f.json $f <<'PY'
import json,sys
l=[x for x in open(sys.argv[1]) if x.startswith("{")]
if l:
    d=json.loads(l[-1]); print(sys.argv[2], f'{d["value"]:.4g}', round(d["ms_per_step"],3), (d.get("cpu_baseline") or {}).get("sample","")[:110])
PY
done
