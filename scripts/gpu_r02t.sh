# N=2 points measured again on the final kernels (shape-class cell kernel, constant face frames)
N=2; tag=r02t; out=gpurun_out; mkdir -p $out
LAUNCH="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N"
run() { name=$1; shift; SECONDS=0; timeout 900 $LAUNCH "$@" --no-cpu-baseline > $out/${tag}_bench_${name}_n$N.json 2> $out/${tag}_bench_${name}_n$N.err; echo "$name exit $? after $SECONDS s"; python - $out/${tag}_bench_${name}_n$N.json <<'PY'
import json,sys
l=[x for x in open(sys.argv[1]) if x.startswith("{")]
if l:
    d=json.loads(l[-1]); r=d["roofline"]
    print("   ", dict(value=f'{d["value"]:.4g}', ms_step=round(d["ms_per_step"],4), face=round(r["ms_face_per_stage"],4), cell=round(r["ms_cell_per_stage"],4), frac=round(r["stage_frac_of_hbm_roofline"],4), e2e=f'{d["e2e"]["value"]:.3g}'))
PY
}
run 64 --steps 20 --warmup 3
run dmach --workload double_mach --scheme modal --limiter eigen --cells 120 --dt 5e-5 --steps 10 --warmup 3
run fstep --workload forward_step --scheme modal --limiter eigen --cells 80 --dt 2e-5 --steps 10 --warmup 3
run 256p3 --cells 256 --degree 3 --steps 5 --warmup 3
