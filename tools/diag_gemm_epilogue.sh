# Diagnostic builds of the GEMM kernel (wrong results on purpose): what is left of a sampler step's GEMM time when the
# output stores, the residual loads and the whole epilogue read-out are removed?  Run on a GPU box; rebuild afterwards.
for flags in "-DB200SD_GEMM_NOSTORE -DB200SD_GEMM_NORES -DB200SD_GEMM_NOEPI" ${DIAG_ALL:+"-DB200SD_GEMM_NOSTORE" "-DB200SD_GEMM_NORES" "-DB200SD_GEMM_NOSTORE -DB200SD_GEMM_NORES"}; do
  B200SD_EXTRA_NVCC_FLAGS="$flags" bash stablediffusion_b200/csrc/build.sh >/dev/null 2>&1
  echo "=== $flags"
  timeout 300 python tools/profile_step.py --batch 8 --out gpurun_out/diag.txt >/dev/null 2>&1; head -1 gpurun_out/diag.txt; grep -E "linear:|conv9:65536x320x2880|conv9:1024x1280x11520" gpurun_out/diag.txt | head -14
done
bash stablediffusion_b200/csrc/build.sh >/dev/null 2>&1
