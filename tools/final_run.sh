python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r02_smoke.log 2>&1
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench.json 2> gpurun_out/r02_bench.err
python bench.py --config 4k --steps 10 --warmup 3 > gpurun_out/r02_bench_4k.json 2> gpurun_out/r02_bench_4k.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r02_bench_ref.json 2> gpurun_out/r02_bench_ref.err
python tools/bench_configs.py > gpurun_out/r02_configs.json 2> gpurun_out/r02_configs.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 3 --lean > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:lk_level_phase_kernel -s 2 -c 1 -o gpurun_out/r02_phase_r7 python bench.py --steps 1 --warmup 3 --lean > gpurun_out/ncu_r7.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:lk_level_phase1_kernel -s 2 -c 1 -o gpurun_out/r02_phase1_r1 python bench.py --steps 1 --warmup 3 --lean --radius 1 > gpurun_out/ncu_r1.log 2>&1
tail -n 2 gpurun_out/r02_smoke.log
