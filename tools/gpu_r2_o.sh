set -x
for C in 512 1024 2048 4096; do bash tools/gpu_variants.sh $C A base 2>&1 | grep VARIANT; done
