python -m pytest tests/test_gpu_diff.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -15
python bench.py --workload diff --no-cpu-baseline --no-extras 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({k:d[k] for k in ('value','e2e','roofline') if k in d}, indent=1))"
