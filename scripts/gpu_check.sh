# quick whole-tree check on one GPU: parity suite, smoke, default bench line
tag=${1:-check}; out=gpurun_out; mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $out/${tag}_pytest_gpu.log; tail -3 $out/${tag}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py --steps 20 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench exit $?"; cut -c1-330 $out/${tag}_bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | cut -c1-200
