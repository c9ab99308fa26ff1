# usage: bash scripts/gpu_r02r.sh N tag  -- the inner | inter split of the cell update on and off, 64^3 FR p=2 on N GPUs
N=$1; tag=$2; out=gpurun_out; mkdir -p $out
LAUNCH="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N"
for sp in 1 0 1 0; do
MCFD_CELL_SPLIT=$sp timeout 600 $LAUNCH --steps 20 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_64_n${N}_split$sp.json 2> $out/${tag}_bench_64_n${N}_split$sp.err; echo "split=$sp exit $?"
python - $out/${tag}_bench_64_n${N}_split$sp.json <<'PY'
import json,sys
l=[x for x in open(sys.argv[1]) if x.startswith("{")]
d=json.loads(l[-1]); r=d["roofline"]
print("   ", dict(value=f'{d["value"]:.4g}', ms_step=round(d["ms_per_step"],4), rep=[round(x,4) for x in d["repeat_ms_per_step"]], face=round(r["ms_face_per_stage"],4), cell=round(r["ms_cell_per_stage"],4), launches=d["gpu_launches"]))
PY
done
