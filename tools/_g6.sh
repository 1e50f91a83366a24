python -m pytest tests -x -q -m gpu 2>&1 | tail -3
for a in "--workload qbb --record --wseed 1" "--workload qbb --record --wseed 0" "--workload qbb --wseed 1" "--workload qq --hidden 16 --prec 0" "--workload qcp --hidden 16 --prec 0"; do python tools/quick.py $a --reps 3 2>&1 | tail -1 | cut -c1-330; done
