nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/fp64_peak minicfd_b200/csrc/tools/fp64_peak.cu && /tmp/fp64_peak > gpurun_out/fp64_peak.json; cat gpurun_out/fp64_peak.json
timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -15
for v in "" _f6 _f8; do timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --lib minicfd_b200/lib/libminicfd_b200$v.so > gpurun_out/bench_v$v.log 2>&1; python - <<PY
import json
l=[x for x in open("gpurun_out/bench_v$v.log") if x.startswith("{")]
d=json.loads(l[-1]) if l else None
print("$v", d and (d["value"], d["ms_per_step"], d["roofline"]["ms_face_per_stage"], d["roofline"]["ms_cell_per_stage"], d["e2e"]["value"]))
PY
done
