for v in b200 d2a d2b d2c; do
  echo "== $v"
  SRE_B200_LIB=simurlacra_b200/lib/libsre_$v.so python bench.py --workload diff --no-cpu-baseline --no-extras 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('fwd+adj', d['value'], 'fwd', d['roofline']['forward']['env_steps_per_s'])"
done
echo "== scalar"
SRE_DIFF_SCALAR=1 python bench.py --workload diff --no-cpu-baseline --no-extras 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('fwd+adj', d['value'], 'fwd', d['roofline']['forward']['env_steps_per_s'])"
