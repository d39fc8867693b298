# ncu evidence for the rows beyond K1-K5 (run under gpurun): two launches of each kernel of scripts/bench_next_rows.py
mkdir -p gpurun_out
CMD="python scripts/bench_next_rows.py"
$CMD > gpurun_out/next_rows_plain.log 2>&1 || exit 1
for k in clip_adam ppo_is_loss n_step_smem sumtree_update sumtree_top fields_copy scan_smem vtrace_fwd_tiled quantile act_epilogue; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 2 -f -o gpurun_out/prof_next_$k $CMD > gpurun_out/ncu_next_$k.log 2>&1
  tail -n 1 gpurun_out/ncu_next_$k.log | cut -c1-120
done
