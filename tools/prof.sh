#!/bin/bash
# usage: tools/prof.sh <tag>  -- ncu --set full of the level-0 strip kernel on the GPU box, then digest here
tag=$1
kern=${2:-lk_level_strip}
shift; shift
/usr/local/graft/bin/gpurun --timeout 600 -- "OF_B200_VARIANT=${OF_B200_VARIANT:-} ncu --set full --clock-control none --import-source on -k regex:$kern -s 2 -c 1 -o gpurun_out/$tag python bench.py --steps 1 --warmup 3 --pairs 16 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; tail -1 gpurun_out/ncu_full.log | cut -c1-80" 2>&1 | tail -2
ncu -i gpurun_out/$tag.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin))
d=dict(zip(rows[0],rows[2]))
for k in ['gpu__time_duration.sum','gcc__cache_requests_type_instruction.sum.pct_of_peak_sustained_elapsed','gcc__cache_requests_type_instruction.sum','sm__icc_request_hit_rate.pct','sm__icc_requests.sum','smsp__inst_executed.sum','smsp__issue_active.avg.pct_of_peak_sustained_active','sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed','dram__bytes_read.sum','dram__bytes_write.sum','launch__registers_per_thread']:
    print('%-90s %s'%(k,d.get(k)))
"
ncu -i gpurun_out/$tag.ncu-rep --page source --csv --print-source sass 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin))
h=rows[1]; ie=h.index('Instructions Executed')
ex=[]
for r in rows[2:]:
    try: ex.append(int(r[ie]))
    except: pass
mx=max(ex)
print('static',len(ex),'max',mx,'hot(>30%max)',sum(1 for e in ex if e>0.3*mx),'warm(3-30%)',sum(1 for e in ex if 0.03*mx<e<=0.3*mx),'cold(<3%,>0)',sum(1 for e in ex if 0<e<=0.03*mx),'never',sum(1 for e in ex if e==0))
"
python tools/ncu_roles.py gpurun_out/$tag.ncu-rep $kern.cu "$@" | grep -E "==|stalls"
