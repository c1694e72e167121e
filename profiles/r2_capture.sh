set -x
for s in 23 22; do GTB200_COUNT_SLICE_LOG2=$s python bench.py --steps 3 --warmup 3 --skip-e2e --no-cpu-baseline > gpurun_out/r2_human_slice$s.json 2>&1; done
timeout 500 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"k_" --csv --log-file gpurun_out/r2_launches_human.csv python bench.py --steps 1 --warmup 1 --skip-e2e --no-cpu-baseline > gpurun_out/ncu_h2.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_tile_sort_keyed|k_keyed_partition_dense|k_keyed_partition_rec|k_kmers" -s 10 -c 4 -o /tmp/r2_prof_human python bench.py --steps 1 --warmup 1 --skip-e2e --no-cpu-baseline > gpurun_out/ncu_h3.log 2>&1
python profiles/summarize.py raw /tmp/r2_prof_human.ncu-rep > gpurun_out/r2_ncu_full_human.txt 2>&1
for k in k_tile_sort_keyed k_keyed_partition_dense k_keyed_partition_rec; do python profiles/summarize.py source /tmp/r2_prof_human.ncu-rep $k 25 >> gpurun_out/r2_ncu_source_human.txt 2>&1; done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_sort_medium|k_stage_partition" -s 3 -c 3 -o /tmp/r2_prof_protein python bench.py --workload protein --steps 1 --warmup 1 --skip-e2e --no-cpu-baseline > gpurun_out/ncu_p.log 2>&1
python profiles/summarize.py raw /tmp/r2_prof_protein.ncu-rep > gpurun_out/r2_ncu_full_protein.txt 2>&1
python profiles/summarize.py source /tmp/r2_prof_protein.ncu-rep k_sort_medium 30 > gpurun_out/r2_ncu_source_protein.txt 2>&1
ls -la gpurun_out /tmp/*.ncu-rep
