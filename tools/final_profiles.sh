# round-end evidence: step profiles (torch.profiler under graph replay), ncu launch list of the default bench step,
# ncu --set full of the HBM / ALU kernels. Run on the GPU box; summaries are made from the files it leaves in gpurun_out/.
mkdir -p gpurun_out/prof
python tools/step_profile.py --graph > gpurun_out/prof/r2_step_profile_nodropout.md 2>/dev/null
python tools/step_profile.py --graph --dropout 0.1 > gpurun_out/prof/r2_step_profile_dropout.md 2>/dev/null
python tools/hbm_ncu.py > gpurun_out/prof/hbm_cases.json 2> gpurun_out/prof/hbm.err && \
ncu --set full --clock-control none --profile-from-start off -f -o gpurun_out/prof/r2b_hbm python tools/hbm_ncu.py > /dev/null 2>> gpurun_out/prof/hbm.err
python bench.py --quick --no-graph --steps 2 --warmup 3 > gpurun_out/prof/quick.json 2> gpurun_out/prof/quick.err && \
ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 1000 -c 700 --csv --log-file gpurun_out/prof/r2b_launches.csv python bench.py --quick --no-graph --steps 2 --warmup 3 > gpurun_out/prof/ncu_quick.log 2>&1
ls -la gpurun_out/prof
