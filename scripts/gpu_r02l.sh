tag=${1:-r02l}; out=gpurun_out; mkdir -p $out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'modal_cell_mma|modal_face_kernel' \
    --launch-skip 8 --launch-count 3 -o $out/${tag}_modal_full -f python bench.py --scheme modal --steps 2 --warmup 3 --no-cpu-baseline \
    > $out/${tag}_ncu_modal_full.log 2>&1; echo "ncu modal full exit $?"
