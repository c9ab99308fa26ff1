# usage: bash scripts/gpu_variants.sh "<bench args A>|<bench args B>|..." [lib suffixes...]  -- bench.py on every (args, build variant) pair
IFS='|' read -ra ARGS <<< "$1"; shift
mkdir -p gpurun_out
for v in "" "$@"; do for a in "${ARGS[@]}"; do
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --lib minicfd_b200/lib/libminicfd_b200$v.so $a > gpurun_out/bench_var.log 2>&1
  python - "$v" "$a" <<'PY'
import json,sys
l=[x for x in open("gpurun_out/bench_var.log") if x.startswith("{")]
d=json.loads(l[-1]) if l else None
print(f"variant '{sys.argv[1]}' args '{sys.argv[2]}'", d and dict(value=f'{d["value"]:.4g}', ms_step=round(d["ms_per_step"],4), ms_face=round(d["roofline"]["ms_face_per_stage"],4), ms_cell=round(d["roofline"]["ms_cell_per_stage"],4)) or open("gpurun_out/bench_var.log").read()[-500:])
PY
done; done
