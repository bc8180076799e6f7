cd $GRAFT_REPO_ROOT
F="python scripts/frame_once.py --workload c3 --frames 4"
python -m pytest tests/test_gpu_rt.py -m gpu -q -x 2>&1 | tail -3
for v in 4 5 6; do $F --variant $v --stats 1 2>&1 | grep -A1 variant; done
for v in 4 5 6; do $F --variant $v 2>&1 | grep variant; done
for t in $((16+8*256)) $((16+16*256)) $((8+16*256)) $((24+12*256)) $((16+24*256)) $((12+4*256)) $((4+4*256)) $((28+28*256)); do echo "tune $t"; $F --variant 6 --tune $t 2>&1 | grep variant; done
for t in $((8+8*256)) $((24+8*256)) $((16+4*256)) $((28+16*256)); do echo "v5 tune $t"; $F --variant 5 --tune $t 2>&1 | grep variant; done
for v in 0 4 6; do python scripts/frame_once.py --workload c5 --variant $v --frames 2 --stats 1 | grep -A1 variant;  python scripts/frame_once.py --workload c5 --variant $v --frames 2 | grep variant; done
