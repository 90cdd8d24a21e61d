set -x
python scripts/dsweep_profile.py c1,c3,c4 device 2 > gpurun_out/r02_dsp.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_dsweep.csv python scripts/dsweep_profile.py c1,c3,c4 device 2 > gpurun_out/ncu1.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:dsweep -c 5 -f -o gpurun_out/r02_dsweep_c4 python scripts/dsweep_profile.py c4 device 1 > gpurun_out/ncu2.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:dsweep -c 5 -f -o gpurun_out/r02_dsweep_c3 python scripts/dsweep_profile.py c3 device 1 > gpurun_out/ncu3.log 2>&1
TCI_DSWEEP_PROF=1 python scripts/dsweep_profile.py c1,c3,c4 device 1 > gpurun_out/r02_dsweep_prof.log 2>&1
python bench.py --steps 2 --warmup 3 --no-sweep > gpurun_out/b_nosweep.log 2>&1 || exit 2
ncu --metrics gpu__time_duration.sum --clock-control none -s 2100 -c 400 --csv --log-file gpurun_out/r02_launches_bench.csv python bench.py --steps 2 --warmup 1 --no-sweep > gpurun_out/ncu4.log 2>&1
ls -la gpurun_out
