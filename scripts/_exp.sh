P=scripts/profile_render.py
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for v in base c4 c6; do
  L=""; [ $v != base ] && L="--lib build_variants/$v.so"
  echo "== $v"; python $P --spp 256 --repeat 2 --profile $L; python $P --spp 256 --repeat 3 $L
done
