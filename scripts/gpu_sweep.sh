# usage: bash scripts/gpu_sweep.sh N tag [tests]   -- the sweeps BASELINE configs[] name, on N GPUs of one box:
#   configs[1] 64^3 FR p=2, configs[2] double Mach (modal p=2 + weno::Eigen, 1.38 M cells), configs[3] forward step
#   (tetrahedra, modal p=2 + weno::Eigen, 1.55 M cells), configs[4] 256^3 FR p=3; with "tests" the multi-GPU parity log first
N=$1; tag=$2; out=gpurun_out; mkdir -p $out
if [ $N -eq 1 ]; then LAUNCH="python bench.py --gpus 1"; else LAUNCH="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N"; fi
if [ "$3" = "tests" ]; then
  timeout 1500 python -m pytest tests/test_multi_gpu.py -m gpu -q > $out/${tag}_pytest_multi_gpu_n$N.log 2>&1
  echo "pytest exit $?" | tee -a $out/${tag}_pytest_multi_gpu_n$N.log; tail -4 $out/${tag}_pytest_multi_gpu_n$N.log
fi
run() { name=$1; shift; SECONDS=0; timeout 900 $LAUNCH "$@" --no-cpu-baseline > $out/${tag}_bench_${name}_n$N.json 2> $out/${tag}_bench_${name}_n$N.err; echo "$name exit $? after $SECONDS s"; tail -n 1 $out/${tag}_bench_${name}_n$N.err; python - $out/${tag}_bench_${name}_n$N.json <<'PY'
import json,sys
l=[x for x in open(sys.argv[1]) if x.startswith("{")]
if l:
    d=json.loads(l[-1]); r=d["roofline"]
    print("   ", dict(value=f'{d["value"]:.4g}', ms_step=round(d["ms_per_step"],4), face=round(r["ms_face_per_stage"],4), cell=round(r["ms_cell_per_stage"],4), frac=round(r["stage_frac_of_hbm_roofline"],4), e2e=f'{d["e2e"]["value"]:.3g}', halo=d["config"].get("halo"), launches=d["gpu_launches"]))
PY
}
run 64 --steps 20 --warmup 3
run dmach --workload double_mach --scheme modal --limiter eigen --cells 120 --dt 5e-5 --steps 10 --warmup 3
run fstep --workload forward_step --scheme modal --limiter eigen --cells 80 --dt 2e-5 --steps 10 --warmup 3
if [ $N -ge 2 ]; then run 256p3 --cells 256 --degree 3 --steps 5 --warmup 3; else echo "256^3 p=3 needs 2 GPUs or more (3 columns x 42.9 GB + flux buffer + tables: out of memory on one, r02j)"; fi
