set -x
cd $GRAFT_REPO_ROOT
python bench.py --steps 40 --warmup 3 > gpurun_out/r02b_bench.json 2> gpurun_out/r02b_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02b_bench_reference.json 2>> gpurun_out/r02b_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02b_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_a.log 2>&1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,dram__throughput.avg.pct_of_peak_sustained_elapsed --clock-control none -k regex:"k_gen_channel|k_gen_rayleigh|k_csi_error|k_prep_realisation|k_prep_los|k_fft4096_rows|k_pa4|k_map_precode|k_propagate" -c 60 --csv --log-file gpurun_out/r02b_hbm_kernels.csv python tools/r02b_hbm_case.py > gpurun_out/ncu_b.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_link_frame_t64 -c 1 -f -o gpurun_out/r02b_t64 python bench.py --steps 1 --warmup 3 --no-cpu --no-secondary --frames 1184 > gpurun_out/ncu_c.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_cnc_loop_t64 -c 1 -f -o gpurun_out/r02b_cnc python tools/r02b_cnc_case.py > gpurun_out/ncu_d.log 2>&1
ls -la gpurun_out
