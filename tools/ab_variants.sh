# usage: tools/ab_variants.sh name1 name2 ...   (under gpurun) -- lib/librpf_b200_<name>.so builds side by side ("base" = the product
# library): quick oracle parity of each variant, then per-box per-kernel milliseconds. BOXES / CONFIG as in tools/perbox.sh.
for v in "$@"; do
  if [ "$v" = base ]; then unset RPF_B200_LIB; else export RPF_B200_LIB=$PWD/raytracer-rpf_b200/lib/librpf_b200_$v.so; fi
  echo "=== variant $v"
  python -m pytest tests/test_gpu_parity.py tests/test_gpu_r2.py -m gpu -q -x -k "box_and_spp_variants or big_fixtures or real" 2>&1 | tail -1
  bash tools/perbox.sh 2>&1 | grep -E "^box|dependency"
done
