P=scripts/profile_render.py
python -m pytest tests -m gpu -x -q -k "bidirectional or bvh" 2>&1 | tail -3
echo "== bdpt"; python $P --spp 32 --repeat 2 --estimator bdpt
echo "== bdpt mesh"; python $P --mesh 8 --size 1024 --spp 16 --repeat 2 --estimator bdpt
