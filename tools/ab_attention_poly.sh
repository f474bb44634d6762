# A/B of the FMA-pipe exp2 share in the softmax (B200SD_ATT_POLY_EVERY: every n-th pair; 0 = all on the MUFU pipe)
for e in ${POLY_SET:-4 0 2 3 8}; do
  B200SD_EXTRA_NVCC_FLAGS="-DB200SD_ATT_POLY_EVERY=$e" bash stablediffusion_b200/csrc/build.sh >/dev/null 2>&1
  echo "=== every $e"
  timeout 300 python tools/bench_ops.py --only ${POLY_ONLY:-attn_self,attn_xl} --out gpurun_out/ab_poly_$e.txt 2>&1 | grep attn
done
bash stablediffusion_b200/csrc/build.sh >/dev/null 2>&1
