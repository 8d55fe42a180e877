#!/bin/bash
# Round-end evidence on one B200: GPU parity suite, both bench arms, the ncu launch list of the bench command, one `--set full`
# capture of the front-end kernels (profile_step = the resident step of the bench workload) and one of the graph-stage kernels.
tag=${1:-final}
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_$tag.log 2>&1; tail -2 gpurun_out/pytest_gpu_$tag.log
python bench.py --impl reference > gpurun_out/bench_ref_$tag.json 2> gpurun_out/bench_ref_$tag.err
python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_$tag.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --e2e-steps 1 --no-file > gpurun_out/ncu_launches_$tag.log 2>&1
ncu --set full --clock-control none --import-source on -k "regex:k_collapse_lut|k_msat|k_select|k_kmer_hash|k_insert_nodes|k_insert_edges|k_list_selected" --launch-skip 7 -c 7 \
    -f -o gpurun_out/full_front_$tag python tools/profile_step.py --scale 1.0 --steps 2 > gpurun_out/ncu_front_$tag.log 2>&1
ncu --set full --clock-control none --import-source on -k 'regex:k_chain_jump|k_onecov_jump|k_onecov_bad|k_thread_steps|k_thread_locate|k_graph_degrees|k_filter_edges' -c 14 \
    -f -o gpurun_out/full_graph_$tag python bench.py --steps 1 --warmup 1 --no-cpu-baseline --e2e-steps 1 --no-file > gpurun_out/ncu_graph_$tag.log 2>&1
tail -c 600 gpurun_out/bench_$tag.json
