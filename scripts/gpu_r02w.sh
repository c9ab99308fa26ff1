tag=r02w; out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests/test_modal.py tests/test_tetra.py tests/test_configs.py tests/test_l1_error.py -m gpu -q > $out/${tag}_pytest.log 2>&1; tail -3 $out/${tag}_pytest.log
run() { name=$1; shift; timeout 600 python bench.py "$@" --no-cpu-baseline > $out/${tag}_bench_$name.json 2> $out/${tag}_bench_$name.err; python - $out/${tag}_bench_$name.json $name <<'PY'
import json,sys
l=[x for x in open(sys.argv[1]) if x.startswith("{")]
d=json.loads(l[-1]); r=d["roofline"]; print(sys.argv[2], dict(value=f'{d["value"]:.4g}', ms_step=round(d["ms_per_step"],4), face=round(r["ms_face_per_stage"],4), cell=round(r["ms_cell_per_stage"],4)))
PY
}
run modal --scheme modal --steps 10 --warmup 3
run modal_eigen --scheme modal --limiter eigen --ic sod --steps 10 --warmup 3
run dmach --workload double_mach --scheme modal --limiter eigen --cells 120 --dt 5e-5 --steps 10 --warmup 3
run fstep --workload forward_step --scheme modal --limiter eigen --cells 80 --dt 2e-5 --steps 10 --warmup 3
