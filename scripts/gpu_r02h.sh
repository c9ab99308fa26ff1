# tuning variants (build-time switches, see minicfd_b200/csrc/Makefile): bench lines only
bash scripts/gpu_variants.sh "--degree 2" _rev _f40 _f36 2>&1 | tee gpurun_out/r02h_variants.txt
