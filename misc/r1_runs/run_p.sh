mkdir -p gpurun_out
python raytrace_b200/build.py --force > /dev/null
for wl in c1 c5 c4; do
  timeout 400 python bench.py --workload $wl --steps 16 --no-cpu-baseline > gpurun_out/r1_p_$wl.json 2> gpurun_out/r1_p_$wl.err; echo "bench $wl rc=$?"; tail -2 gpurun_out/r1_p_$wl.err
done
python - <<'PY'
import json
for wl in ("c1","c5","c4"):
    try:
        d=json.load(open(f"gpurun_out/r1_p_{wl}.json")); print(wl, "Mrays/s %.0f ms %.4f fps %.0f e2e fps %.0f"%(d["value"], d["ms_per_step"], d["frames_per_s"], d["e2e"]["frames_per_s"]), d["kernel_ms_per_step_rank0"], "steps/ray %.1f"%d["mean_steps_per_ray"], "limit", d["limit_exhausted"])
    except Exception as e: print(wl, "failed", e)
PY
bash misc/sweep.sh q_u1 "" q_u2 "-DVXRT_DN_UNROLL=2" q_u4 "-DVXRT_DN_UNROLL=4" q_u6 "-DVXRT_DN_UNROLL=6"
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/sweep_c2_q_*.json")):
    try:
        d=json.load(open(f)); print(f, d["kernel_ms_per_step_rank0"])
    except Exception as e: print(f, e)
PY
