python -m pytest tests -x -q -m gpu 2>&1 | tail -3
echo "== bulk"; python tools/quick.py --workload qbb --record --wseed 1 --reps 3 2>&1 | tail -1 | cut -c1-330
echo "== no bulk"; SRE_NO_BULK=1 python tools/quick.py --workload qbb --record --wseed 1 --reps 3 2>&1 | tail -1 | cut -c1-330
echo "== fb3 bulk"; SRE_B200_LIB=simurlacra_b200/lib/libsre_fb3.so python tools/quick.py --workload qbb --record --wseed 1 --reps 3 2>&1 | tail -1 | cut -c1-330
echo "== fb3 norecord"; SRE_B200_LIB=simurlacra_b200/lib/libsre_fb3.so python tools/quick.py --workload qbb --wseed 1 --reps 3 2>&1 | tail -1 | cut -c1-330
echo "== bulk wseed0"; python tools/quick.py --workload qbb --record --wseed 0 --reps 3 2>&1 | tail -1 | cut -c1-330
echo "== qq records TC"; python tools/quick.py --workload qq --record --reps 3 2>&1 | tail -1 | cut -c1-330
echo "== qq records TC nobulk"; SRE_NO_BULK=1 python tools/quick.py --workload qq --record --reps 3 2>&1 | tail -1 | cut -c1-330
