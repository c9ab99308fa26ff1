out=gpurun_out; tag=r02z
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'modal_' --launch-skip 8 --launch-count 7 -o $out/${tag}_modal_full -f \
  python bench.py --scheme modal --limiter eigen --ic sod --steps 2 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_modal_full.log 2>&1; echo "ncu exit $?"
timeout 300 python bench.py --scheme modal --limiter eigen --ic sod --steps 10 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_modal_eigen.json 2>/dev/null; cut -c1-200 $out/${tag}_bench_modal_eigen.json
