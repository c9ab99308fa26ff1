bash scripts/gpu_variants.sh "--scheme modal" _w2 _w8 _w4b5 2>&1 | tee gpurun_out/r02v_variants.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'modal_cell_mma' \
    --launch-skip 4 --launch-count 1 -o gpurun_out/r02v_mma_full -f python bench.py --scheme modal --steps 2 --warmup 3 --no-cpu-baseline \
    > gpurun_out/r02v_ncu.log 2>&1; echo "ncu exit $?"
