tag=${1:-r02n}; out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests/test_modal.py tests/test_tetra.py tests/test_configs.py -m gpu -q -x > $out/${tag}_pytest.log 2>&1; grep -v "^  File\|^$" $out/${tag}_pytest.log | tail -8
bash scripts/gpu_sweep.sh 1 $tag 2>&1 | grep -v 256p3 | head -12
