# quick GPU check of a kernel change: parity subset (skipped with NOTEST=1), then one bench line per variant given as
# "ENV=.. ENV2=.. [LIBSUF=_x]" (LIBSUF picks minicfd_b200/lib/libminicfd_b200_x.so)
out=gpurun_out; mkdir -p $out
if [ -z "$NOTEST" ]; then timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_golden.py tests/test_size_parity.py tests/test_configs.py -m gpu -q -x 2>&1 | tail -4; fi
for v in "$@"; do
  echo "== $v"
  lib=""; for kv in $v; do case $kv in LIBSUF=*) lib="--lib minicfd_b200/lib/libminicfd_b200${kv#LIBSUF=}.so";; esac; done
  env $v timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline $lib $BENCHARGS 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print(dict(ms_step=round(d['ms_per_step'],4), med=round(d.get('median_ms_per_step',0),4), face=round(r['ms_face_per_stage'],4), cell=round(r['ms_cell_per_stage'],4), frac=round(r['stage_frac_of_hbm_roofline'],4), finite=d['solution_finite']))
    elif 'rror' in l: print(l.rstrip())
"
done
