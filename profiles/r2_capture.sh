# Round-2 evidence run (one B200): bench lines of every workload, the reference arm, the ncu launch list and one
# `--set full` capture of the hot kernels at 3.1 Gbp.  Large .ncu-rep files stay on the box; only summaries come back.
set -x
python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_human_1gpu.json 2> gpurun_out/r2_bench_human_1gpu.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_bench_human_reference.json 2>&1
python bench.py --workload chr1 --steps 10 --warmup 3 > gpurun_out/r2_bench_chr1_1gpu.json 2>&1
python bench.py --workload protein --steps 5 --warmup 3 > gpurun_out/r2_bench_protein_1gpu.json 2>&1
python bench.py --workload ecoli --steps 20 --warmup 5 > gpurun_out/r2_bench_ecoli_1gpu.json 2>&1
python bench.py --workload genomelike --steps 5 --warmup 3 > gpurun_out/r2_bench_genomelike_1gpu.json 2>&1
timeout 500 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"k_" --csv --log-file gpurun_out/r2_launches_human.csv python bench.py --steps 1 --warmup 1 --skip-e2e --no-cpu-baseline > gpurun_out/ncu_h2.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_tile_sort_keyed|k_keyed_partition_dense|k_keyed_partition_rec|k_count_dna" -s 7 -c 8 -o /tmp/r2_prof_human python bench.py --steps 1 --warmup 1 --skip-e2e --no-cpu-baseline > gpurun_out/ncu_h3.log 2>&1
python profiles/summarize.py raw /tmp/r2_prof_human.ncu-rep > gpurun_out/r2_ncu_full_human.txt 2>&1
rm -f gpurun_out/r2_ncu_source_human.txt
for k in k_tile_sort_keyed k_keyed_partition_dense k_keyed_partition_rec k_count_dna; do python profiles/summarize.py source /tmp/r2_prof_human.ncu-rep $k 25 >> gpurun_out/r2_ncu_source_human.txt 2>&1; done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_tile_subsort|k_keyed_partition_bytes|k_keyed_partition_rec" -s 3 -c 3 -o /tmp/r2_prof_protein python bench.py --workload protein --steps 1 --warmup 1 --skip-e2e --no-cpu-baseline > gpurun_out/ncu_p.log 2>&1
python profiles/summarize.py raw /tmp/r2_prof_protein.ncu-rep > gpurun_out/r2_ncu_full_protein_keyed.txt 2>&1
python profiles/summarize.py source /tmp/r2_prof_protein.ncu-rep k_tile_subsort 25 > gpurun_out/r2_ncu_source_protein_keyed.txt 2>&1
ls -la gpurun_out /tmp/*.ncu-rep
