# The FMA experiment of SURVEY.md section 7 / VERDICT item 5(ii): a second library with dot4's
# products fused (PT_FMA=1), measured next to the shipped one and put through the gates
# (scaler counts integer-identical, lnL within 1e-9 of the reference at the full size, walk tables).
# Build here (CPU box): bash scripts/fma_variant.sh build ; run under gpurun: bash scripts/fma_variant.sh run
set -e
if [ "$1" = build ]; then
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-ffp-contract=off \
    -shared -DPT_FMA=1 -o pt_b200/lib/libptgpu_fma.so pt_b200/csrc/engine.cu pt_b200/csrc/gtr_model.cpp
  exit 0
fi
export PTGPU_LIB=$PWD/pt_b200/lib/libptgpu_fma.so
python bench.py --steps 5 --warmup 3 --no-sharded --no-cpu-baseline > gpurun_out/r2_fma_bench.json 2> gpurun_out/r2_fma_bench.err || tail -3 gpurun_out/r2_fma_bench.err
python -m pytest tests/test_gpu_configs.py tests/test_gpu_parity.py -m gpu -q > gpurun_out/r2_fma_tests.log 2>&1 || true
tail -15 gpurun_out/r2_fma_tests.log | cut -c1-200
