# star_u member + tetrahedra: GPU parity of the whole tree, then the default bench line (face-kernel cost of the extra store)
tag=${1:-r02i}; out=gpurun_out; mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=5 > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $out/${tag}_pytest_gpu.log
tail -12 $out/${tag}_pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench exit $?"; cut -c1-600 $out/${tag}_bench.json
timeout 400 python bench.py --scheme modal --steps 10 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_modal.json 2> $out/${tag}_bench_modal.err; echo "modal exit $?"; cut -c1-300 $out/${tag}_bench_modal.json
