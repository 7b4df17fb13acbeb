#!/bin/bash
# Turns the .ncu-rep files / launch list of tools/r2_final.sh (gpurun_out/<tag>_*) into the text summaries under profiles/.
# Runs on the CPU box (needs the ncu CLI, cuobjdump, nvdisasm).   usage: tools/make_profiles.sh <tag>
tag=${1:-r02c}
G=gpurun_out; P=profiles; T=/tmp/prof_$tag
rm -rf $T; mkdir -p $T; ( cd $T && cuobjdump -xelf all $OLDPWD/ray-tracing-one-weekend_b200/csrc/librtow_b200.so > /dev/null && for f in *.cubin; do nvdisasm -g -c $f > ${f%.cubin}.sass 2>/dev/null; done )
lines() { ncu -i $G/$1.ncu-rep --page source --csv > $T/$1.csv 2>/dev/null; python tools/ncu_lines.py $T/$1.csv $T/$2.sm_100a.sass $3 ${4:-30} ${5:-1}; }
{
  echo "ncu --set full --clock-control none --import-source on -k regex:<kernel> -s 60 -c 1 python tools/prof_render.py --scene random_ned --spp 128   ($tag; configs[1] scene with its 2048x1024 image texture, 1920x1080, full 64-sample blocks, stream of 2^23 entries of 64 B)"
  python tools/ncu_summary.py $G/${tag}_extend.ncu-rep
  echo; echo "---- wf_extend_kernel: source lines (ncu SASS page joined with nvdisasm line info, tools/ncu_lines.py)"
  lines ${tag}_extend wavefront wf_extend_kernelILi8 30
  echo; python tools/ncu_summary.py $G/${tag}_shade.ncu-rep
  echo; echo "---- wf_shade_kernel: source lines"
  lines ${tag}_shade wavefront wf_shade_kernel 30
} > $P/${tag}_wavefront_ncu_summary.txt 2>&1
{
  echo "the same captures on the Cornell + smoke scene (cornell_smoke_ned, 1920x1080, 128 spp): the list-scene extend kernel and the shade kernel"
  python tools/ncu_summary.py $G/${tag}_list.ncu-rep
  echo; echo "---- wf_extend_list_kernel: source lines"
  lines ${tag}_list wavefront wf_extend_list_kernel 30
  echo; python tools/ncu_summary.py $G/${tag}_shade_smoke.ncu-rep
} > $P/${tag}_list_scene_ncu_summary.txt 2>&1
{
  echo "ncu --set full --clock-control none --import-source on -k regex:trace_persistent_kernel -c 4 python tools/prof_trace.py --rays 22   ($tag; config 5: 1 M spheres, 2^22 rays per class; launches: coherent warm-up, coherent timed, coherent counting build (2^16 rays), incoherent)"
  python tools/ncu_summary.py $G/${tag}_trace.ncu-rep
  echo; echo "---- trace_persistent_kernel, coherent set: source lines"
  lines ${tag}_trace kernels trace_persistent_kernelILb0ELi5 25 2
  echo; echo "---- trace_persistent_kernel, incoherent set: source lines"
  lines ${tag}_trace kernels trace_persistent_kernelILb0ELi5 25 4
} > $P/${tag}_trace_ncu_summary.txt 2>&1
head -1500 $G/${tag}_launches_bench.csv > $P/${tag}_launches_bench.csv
python - $G/${tag}_launches_bench.csv $G/${tag}_plain.json > $P/${tag}_launch_shares.txt <<'PY'
import csv, json, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr = [r for r in rows if "Kernel Name" in r][0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
tot = collections.Counter(); cnt = collections.Counter()
for r in rows:
    if len(r) == len(hdr) and r[0].isdigit():
        k = r[ki].split("(")[0].replace("void ", "").replace("rtow::", "")
        tot[k] += float(r[vi].replace(",", "")); cnt[k] += 1
allns = sum(tot.values())
print("ncu launch list of `python bench.py --steps 1 --warmup 3 --spp 128 --no-cpu --no-bvh --no-extra` (gpu__time_duration.sum per launch, --clock-control none; first 3000 launches; cold-cache and serialised: compare SHARES, not absolutes)")
for k, v in tot.most_common():
    print(f"  {k:42s} {cnt[k]:5d} launches  {v / 1e6:9.3f} ms  {100 * v / allns:5.1f} %   avg {v / cnt[k] / 1e3:8.1f} us")
d = json.load(open(sys.argv[2])); r = d["roofline"]
e, s = r["kernel_ms"], r["shade_kernel_ms"]
print(f"bench.py's own CUDA-event timing of the same command, without ncu: extend {e * 1e3:.1f} us, shade {s * 1e3:.1f} us per full-stream iteration -> extend share {100 * e / (e + s):.1f} % of the pair")
ext = sum(v for k, v in tot.items() if "wf_extend" in k); sh = tot.get("wf_shade_kernel", 0.0)
print(f"ncu: extend share of the pair {100 * ext / (ext + sh):.1f} %  (library kernels in the list: {sum(1 for k in tot if not k.startswith(('wf_', 'resolve', 'trace', 'render')))} kinds: {[k for k in tot if not k.startswith(('wf_', 'resolve', 'trace', 'render'))]})")
PY
python - <<PY
import json, subprocess, csv, io
def dram(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--metrics", "dram__bytes_read.sum,dram__bytes_write.sum"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h = rows[0]; r = rows[2]
    conv = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    units = rows[1]
    rd = float(r[h.index("dram__bytes_read.sum")]) * conv[units[h.index("dram__bytes_read.sum")]]
    wr = float(r[h.index("dram__bytes_write.sum")]) * conv[units[h.index("dram__bytes_write.sum")]]
    return int(rd), int(wr)
t = {"_comment": "dram__bytes_read.sum + dram__bytes_write.sum per launch from the ncu --set full --clock-control none captures summarised in profiles/${tag}_*_ncu_summary.txt (stream of 2^23 entries of 64 B). bench.py copies the dominant kernel's figure into roofline.traffic when it runs the same scene and stream size."}
for k, rep in (("wf_extend_kernel", "$G/${tag}_extend.ncu-rep"), ("wf_shade_kernel", "$G/${tag}_shade.ncu-rep"), ("wf_extend_list_kernel", "$G/${tag}_list.ncu-rep")):
    rd, wr = dram(rep)
    t[k] = {"pool_slots": 8388608, "dram_bytes_per_launch": rd + wr, "read": rd, "written": wr}
json.dump(t, open("$P/ncu_traffic.json", "w"), indent=2)
print(json.dumps(t, indent=1)[:600])
PY
ls -la $P/${tag}_*
