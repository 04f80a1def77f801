P=scripts/profile_render.py
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
echo "== base"; python $P --spp 256 --repeat 2 --profile; python $P --spp 256 --repeat 3
echo "== lt"; python $P --spp 256 --repeat 2 --estimator lt
echo "== bdpt"; python $P --spp 32 --repeat 2 --estimator bdpt
