tag=${1:-r02k}; out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests/test_modal.py tests/test_tetra.py tests/test_configs.py -m gpu -q -x > $out/${tag}_pytest.log 2>&1; grep -v "^  File\|^$" $out/${tag}_pytest.log | tail -30
for tb in 1 0; do
MCFD_MODAL_TABLES=$tb timeout 400 python bench.py --scheme modal --steps 10 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_modal_t$tb.json 2> $out/${tag}_bench_modal_t$tb.err; echo "modal tables=$tb exit $?"
python - $out/${tag}_bench_modal_t$tb.json <<'PY'
import json,sys
l=[x for x in open(sys.argv[1]) if x.startswith("{")]
d=json.loads(l[-1]); r=d["roofline"]; print(dict(value=f'{d["value"]:.4g}', ms_step=round(d["ms_per_step"],4), face=round(r["ms_face_per_stage"],4), cell=round(r["ms_cell_per_stage"],4)))
PY
done
