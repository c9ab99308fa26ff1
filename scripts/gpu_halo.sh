# usage (GPU box with N >= 2 GPUs): bash scripts/gpu_halo.sh <tag> <N>
# Multi-GPU parity (both halo transports) and the strong-scaling bench line at N GPUs with each transport.
tag=${1:-halo}; n=${2:-2}
out=gpurun_out
mkdir -p $out
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -4 $out/${tag}_pytest.log
for t in 1 0; do
  MCFD_HALO_P2P=$t timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 \
    --master-port $((29511 + t)) bench.py --gpus $n --steps 20 --warmup 3 --no-cpu-baseline \
    > $out/${tag}_bench_n${n}_p2p$t.json 2> $out/${tag}_bench_n${n}_p2p$t.err; echo "bench N=$n p2p=$t exit $?"
  tail -c 1200 $out/${tag}_bench_n${n}_p2p$t.json | tr ',' '\n' | grep -E 'ms_per_step|"value"|halo|n_gpus' | head -6
done
