"""profiles/r02_* from the files tests/gpu_r02_final.sh leaves in gpurun_out/r02f: ncu details / per-line / SASS summaries and
key metrics of the three captures, traffic.json (stamped), bench lines, launch list, test tail.
    python tools/profiles_from_campaign.py [gpurun_out/r02f]"""
import csv
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
O = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "r02f")
P = os.path.join(ROOT, "profiles")
CAPS = (("prof", "render_frame", "render.map 3840x2160, history frame"),
        ("prof_wide", "render_frame_wide", "1,024-sphere grid 3840x2160"),
        ("prof_tree", "render_tree", "one N=8 part of render.map 3840x2160, history frame"),
        ("prof_tacc", "render_frame_tri_accel", "49,928-triangle mesh 3840x2160 through the N4 box hierarchy (rt_kernels_tri_accel.cu), history frame"))
names = [l.split()[0] for l in open(os.path.join(P, "r01_v11_render_frame_key_metrics.txt")).read().split("\n") if l and not l.startswith("Kernel")]
names += ["smsp__inst_executed.sum", "sm__inst_issued.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
          "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__grid_size"]


def ncu(rep, *args):
    return subprocess.run(["ncu", "-i", rep] + list(args), capture_output=True, text=True).stdout


for n, t, what in CAPS:
    rep = os.path.join(O, n + ".ncu-rep")
    if not os.path.exists(rep):
        print("missing", rep)
        continue
    open(os.path.join(P, "r02_%s_details.txt" % t), "w").write(ncu(rep, "--page", "details"))
    src = "/tmp/%s.src.csv" % n
    open(src, "w").write(ncu(rep, "--page", "source", "--csv", "--print-source", "cuda,sass"))
    open(os.path.join(P, "r02_%s_lines.txt" % t), "w").write(
        subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_line_summary.py"), src, "40"], capture_output=True, text=True).stdout)
    sass = "/tmp/%s.sass.csv" % n
    open(sass, "w").write(ncu(rep, "--page", "source", "--csv", "--print-source", "sass"))
    open(os.path.join(P, "r02_%s_sass_summary.txt" % t), "w").write(
        subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_sass_summary.py"), sass], capture_output=True, text=True).stdout)
    rows = list(csv.reader(ncu(rep, "--page", "raw", "--csv").splitlines()))
    h, u, v = rows[0], rows[1], rows[2]
    d = {a: (b, c) for a, b, c in zip(h, u, v)}
    with open(os.path.join(P, "r02_%s_key_metrics.txt" % t), "w") as f:
        f.write("Kernel Name  %s   (%s)\n" % (d["Kernel Name"][1], what))
        for k in dict.fromkeys(names):
            if k in d:
                f.write("%s %s %s\n" % (k, d[k][0], d[k][1]))
    print(t, d["Kernel Name"][1][:50], d["gpu__time_duration.sum"], d.get("launch__grid_size"))
subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_traffic.py"), "render_map_4k=" + os.path.join(O, "prof.ncu-rep"),
                "spheres1024_4k=" + os.path.join(O, "prof_wide.ncu-rep")], check=True)
for src, dst in (("bench_n1.json", "r02_final_bench_n1.json"), ("bench_ref.json", "r02_final_bench_ref.json"), ("launches.csv", "r02_final_launches.csv")):
    shutil.copy(os.path.join(O, src), os.path.join(P, dst))
open(os.path.join(P, "r02_final_pytest_gpu_and_smoke.txt"), "w").write(open(os.path.join(O, "pytest_gpu.txt")).read() + open(os.path.join(O, "smoke.txt")).read())
