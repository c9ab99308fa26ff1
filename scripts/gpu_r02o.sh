tag=${1:-r02o}; out=gpurun_out; mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $out/${tag}_pytest_gpu.log; tail -4 $out/${tag}_pytest_gpu.log
bash scripts/gpu_sweep.sh 1 $tag 2>&1 | head -12
timeout 600 python bench.py --n 128 --degree 3 --steps 3 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_128_p3.json 2> $out/${tag}_bench_128_p3.err; echo "128^3 p3 exit $?"; cut -c1-200 $out/${tag}_bench_128_p3.json
