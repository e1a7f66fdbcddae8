# usage: PATTERN="membership|ordered" tools/ab_kernels.sh name1 name2 ...   (under gpurun) -- per-box milliseconds of the kernels matching
# PATTERN for each lib/librpf_b200_<name>.so ("base" = the product library). BOXES / CONFIG as in tools/perbox.sh.
for v in "$@"; do
  if [ "$v" = base ]; then unset RPF_B200_LIB; else export RPF_B200_LIB=$PWD/raytracer-rpf_b200/lib/librpf_b200_$v.so; fi
  echo "=== variant $v"
  bash tools/perbox.sh 2>&1 | grep -E "^box|${PATTERN:-membership|ordered}"
done
