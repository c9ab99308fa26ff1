# usage: bash scripts/gpu_multi.sh N tag [tests-k-expression|none]   -- multi-GPU parity log + the sweeps configs[] name, on N GPUs
N=$1; tag=$2; kexpr=$3; out=gpurun_out; mkdir -p $out
if [ $N -eq 1 ]; then LAUNCH="python bench.py --gpus 1"; else LAUNCH="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N"; fi
if [ "$kexpr" != "none" ]; then
  if [ -n "$kexpr" ]; then timeout 1500 python -m pytest tests/test_multi_gpu.py -m gpu -q -k "$kexpr" > $out/${tag}_pytest_multi_gpu_n$N.log 2>&1
  else timeout 1500 python -m pytest tests/test_multi_gpu.py -m gpu -q > $out/${tag}_pytest_multi_gpu_n$N.log 2>&1; fi
  echo "pytest exit $?" | tee -a $out/${tag}_pytest_multi_gpu_n$N.log; tail -4 $out/${tag}_pytest_multi_gpu_n$N.log
fi
timeout 600 $LAUNCH --steps 20 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_64_n$N.json 2> $out/${tag}_bench_64_n$N.err; echo "64^3 exit $?"; cut -c1-300 $out/${tag}_bench_64_n$N.json
timeout 900 $LAUNCH --workload double_mach --scheme modal --limiter eigen --cells 120 --dt 5e-5 --steps 10 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_dmach_n$N.json 2> $out/${tag}_bench_dmach_n$N.err; echo "double mach exit $?"; cut -c1-300 $out/${tag}_bench_dmach_n$N.json
if [ $N -ge 2 ]; then
  timeout 1500 $LAUNCH --cells 256 --degree 3 --steps 5 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_256p3_n$N.json 2> $out/${tag}_bench_256p3_n$N.err; echo "256^3 p3 exit $?"; cut -c1-300 $out/${tag}_bench_256p3_n$N.json
fi
tail -n 3 $out/${tag}_bench_*_n$N.err
