tag=${1:-r02b}; out=gpurun_out; mkdir -p $out
timeout 2400 python -m pytest tests -m gpu -q --durations=10 > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $out/${tag}_pytest_gpu.log
tail -25 $out/${tag}_pytest_gpu.log
( time timeout 900 python bench.py --impl reference --steps 3 --warmup 1 ) > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err; echo "ref exit $?"; tail -c 1500 $out/${tag}_bench_ref.json; tail -4 $out/${tag}_bench_ref.err
timeout 600 python bench.py --steps 20 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench exit $?"; tail -c 1800 $out/${tag}_bench.json
