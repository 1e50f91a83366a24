set -u
mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/r02_pytest.log)
tail -3 gpurun_out/r02_pytest.log
timeout 600 python bench.py > gpurun_out/r02_bench.json 2> gpurun_out/r02_bench.err; echo "bench rc $?"
timeout 600 python bench.py --impl reference > gpurun_out/r02_bench_ref.json 2> gpurun_out/r02_bench_ref.err; echo "ref rc $?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/r02_ncu_launch.log 2>&1; echo "launchlist rc $?"
timeout 600 bash tools/prof.sh r02_a_headline --workload qq --prec 1
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')"
