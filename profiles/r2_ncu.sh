#!/bin/bash
# Round 2 ncu evidence (one GPU).  Every ncu pass follows a plain run of the same command that exited 0.
out=gpurun_out; mkdir -p $out; tag=${1:-r2n}
EXTRA=smsp__inst_executed_op_global_red.sum,smsp__inst_executed_op_global_atom.sum,l1tex__t_set_accesses_pipe_lsu_mem_global_op_red.sum,l1tex__t_set_accesses_pipe_lsu_mem_global_op_atom.sum,l1tex__t_requests_pipe_lsu_mem_global_op_red.sum,l1tex__t_requests_pipe_lsu_mem_global_op_atom.sum,lts__t_requests_op_red.sum,lts__t_requests_op_atom.sum,lts__t_sectors_op_red.sum,lts__t_sectors_op_atom.sum,smsp__inst_executed_pipe_lsu.sum,smsp__inst_executed_pipe_fp64.sum
for v in grad base; do
  python profiles/run_pipeline.py $v > $out/${tag}_plain_$v.log 2>&1 &&
  timeout 600 ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:track_kernel -s 2 -c 1 -f -o $out/${tag}_prof_track_$v \
      python profiles/run_pipeline.py $v > $out/${tag}_ncu_track_$v.log 2>&1
  tail -n 2 $out/${tag}_ncu_track_$v.log
done
python profiles/run_pipeline.py grad > $out/${tag}_plain_grad2.log 2>&1 &&
timeout 600 ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:normals_compact -s 6 -c 1 -f -o $out/${tag}_prof_normals \
    python profiles/run_pipeline.py grad > $out/${tag}_ncu_normals.log 2>&1
tail -n 2 $out/${tag}_ncu_normals.log
python bench.py --steps 12 --warmup 3 --no-extra --no-room --reps 1 --cpu-budget-s 0.5 > $out/${tag}_bench_steps12.json 2>/dev/null &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file $out/${tag}_launches_bench_steps12.csv \
    python bench.py --steps 12 --warmup 3 --no-extra --no-room --reps 1 --cpu-budget-s 0.5 > $out/${tag}_ncu_launches.log 2>&1
python profiles/launch_summary.py $out/${tag}_launches_bench_steps12.csv > $out/${tag}_launch_summary.txt 2>&1; cat $out/${tag}_launch_summary.txt
ls -la $out | grep ${tag}
