O=gpurun_out
python -m pytest tests -m gpu -x -q > $O/f_pytest.log 2>&1; tail -2 $O/f_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/f_smoke.log 2>&1; tail -1 $O/f_smoke.log
python bench.py --impl reference --steps 2 --warmup 1 > $O/f_bench_ref.json 2> $O/f_bench_ref.err
python bench.py > $O/f_bench.json 2> $O/f_bench.err
cut -c1-200 $O/f_bench.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/f_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/f_ncu_bench.log 2>&1
P=scripts/profile_render.py
( echo "== lt cornell 1024 512spp"; python $P --size 1024 --spp 512 --estimator lt --repeat 2
  echo "== bdpt cornell 1024 64spp"; python $P --size 1024 --spp 64 --estimator bdpt --repeat 2
  echo "== pt mesh8 2048 16spp"; python $P --mesh 8 --size 2048 --spp 16 --repeat 2 --profile
  echo "== lt mesh8 2048 16spp"; python $P --mesh 8 --size 2048 --spp 16 --estimator lt --repeat 2
  echo "== bdpt mesh8 1024 16spp"; python $P --mesh 8 --size 1024 --spp 16 --estimator bdpt --repeat 2 ) > $O/f_estimators.log 2>&1
cat $O/f_estimators.log
