set -x
cd $GRAFT_REPO_ROOT
F="python scripts/frame_once.py --workload c3 --frames 4"
for t in 0 12 16 24 28; do $F --variant 4 --tune $t; done 2>&1 | grep variant
for e in st8 mb3 st8mb5; do SKYBOX_RT_LIB=skybox_rt_b200/lib/exp_$e/libvortex.so $F --variant 4 2>&1 | grep variant | sed "s/^/$e /"; done
$F --variant 5; $F --variant 0
python scripts/frame_once.py --workload c5 --variant 4 --frames 2
python scripts/frame_once.py --workload c5 --variant 0 --frames 2
python scripts/frame_once.py --workload c2 --variant 4 --frames 3
python scripts/frame_once.py --workload c2 --variant 0 --frames 3
python -c "
from skybox_rt_b200 import vortex as vx
with vx.Device() as d:
    for mb in (16,32,64,96,256,1024,4096):
        print('probe', mb, 'MiB', round(d.read_bandwidth_gbs(mb<<20, 20),1), 'GB/s')
"
$F --variant 4 > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_wave_launches.csv $F --variant 4 > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:rt_wave_trace -s 9 -c 3 -o gpurun_out/r02_wave_trace $F --variant 4 > gpurun_out/ncu2.log 2>&1
