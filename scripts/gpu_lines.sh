# usage (GPU box): bash scripts/gpu_lines.sh  -- the two cell kernels (node-per-thread / line-per-thread, MCFD_CELL_LINES=0|1)
# at every degree: parity of each, then the bench line of each
mkdir -p gpurun_out
for l in 1 0; do
  MCFD_CELL_LINES=$l timeout 400 python -m pytest tests/test_gpu_parity.py tests/test_golden.py tests/test_configs.py tests/test_dropin.py -m gpu -x -q 2>&1 | tail -1
done
run() {
  MCFD_CELL_LINES=$1 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline $2 > gpurun_out/bench_var.log 2>&1
  python - "$1" "$2" <<'PY'
import json,sys
l=[x for x in open("gpurun_out/bench_var.log") if x.startswith("{")]
d=json.loads(l[-1]) if l else None
print(f"lines={sys.argv[1]} args '{sys.argv[2]}'", d and dict(value=f'{d["value"]:.4g}', ms_step=round(d["ms_per_step"],4), ms_face=round(d["roofline"]["ms_face_per_stage"],4), ms_cell=round(d["roofline"]["ms_cell_per_stage"],4)) or open("gpurun_out/bench_var.log").read()[-500:])
PY
}
for l in 0 1; do run $l ""; run $l "--scheme dgsem"; run $l "--degree 1"; run $l "--degree 3 --n 48"; done
