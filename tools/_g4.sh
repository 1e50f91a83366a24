python -m pytest tests/test_gpu_diff.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -3
python bench.py --workload diff --no-cpu-baseline --no-extras 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps(d['roofline'], indent=1))"
bash tools/prof.sh r02_c_c4 --workload qbb --record --envs 1048576 --wseed 1
