N=$1; out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -q -k "halo_exchange or ragged" > $out/pack_pytest_n$N.log 2>&1; echo "pytest exit $?"; tail -3 $out/pack_pytest_n$N.log
LAUNCH="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N"
timeout 600 $LAUNCH --steps 20 --warmup 3 --no-cpu-baseline > $out/r02z_bench_64_n$N.json 2> $out/r02z_bench_64_n$N.err; echo "bench exit $?"
python - $out/r02z_bench_64_n$N.json <<'PY'
import json,sys
l=[x for x in open(sys.argv[1]) if x.startswith("{")]
d=json.loads(l[-1]); r=d["roofline"]
print("   ", dict(value=f'{d["value"]:.4g}', ms_step=round(d["ms_per_step"],4), rep=[round(x,4) for x in d["repeat_ms_per_step"]], face=round(r["ms_face_per_stage"],4), cell=round(r["ms_cell_per_stage"],4)))
PY
