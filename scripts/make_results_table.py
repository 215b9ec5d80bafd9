#!/usr/bin/env python
"""Builds BASELINE.md §5 ("Results") from the committed bench lines under profiles/ (r02z_*).

  python scripts/make_results_table.py            # prints the markdown
  python scripts/make_results_table.py --write    # replaces §5 of BASELINE.md
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = os.path.join(ROOT, "profiles")
PEAK_NOMINAL = 8000.0


def load_line(name):
    path = os.path.join(P, name)
    if not os.path.exists(path):
        return None
    lines = [l for l in open(path) if l.startswith("{")]
    return json.loads(lines[-1]) if lines else None


def load_lines(name):
    path = os.path.join(P, name)
    return [json.loads(l) for l in open(path) if l.startswith("{")] if os.path.exists(path) else []


def codec_records(line):
    recs = {"lz4": line}
    recs.update(line.get("codecs", {}))
    return recs


def main():
    out = []
    b1 = load_line("r02z_bench_1gpu.json")
    ref1 = load_line("r02z_reference_1gpu.json")
    refgpu = {l["metric"].split("_")[0]: l for l in load_lines("r02z_refgpu.jsonl")}
    peak = b1["roofline"]["peak"]
    out.append(f"Round 2, one B200 (measured HBM peak {peak:.0f} GB/s, `MEASURED_PEAKS.json`; nominal {PEAK_NOMINAL:.0f}).  Source lines: "
               "`profiles/r02z_bench_1gpu.json` (this library, `python bench.py`), `profiles/r02z_reference_1gpu.json` "
               "(`--impl reference`, CPU arm on the full config), `profiles/r02z_refgpu.jsonl` (`--impl refgpu`, the reference's own GPU build).\n")
    out.append("### 5a. Headline configs, N = 1 (GB/s of uncompressed data)\n")
    out.append("| config | codec | direction | GB/s | ratio | algorithmic GB/s | % of measured peak | % of 8000 | reference GPU GB/s | CPU GB/s (all cores) | CPU GB/s (1 core) |")
    out.append("|---|---|---|---|---|---|---|---|---|---|---|")
    names = {"lz4": "C2", "snappy": "C1", "cascaded": "C3"}
    for codec, r in codec_records(b1).items():
        cpu = r.get("cpu_baseline", {})
        st = cpu.get("single_thread", {})
        g = refgpu.get(codec, {})
        for d in ("compress", "decompress"):
            roof = r[f"roofline_{d}"]
            out.append(f"| {names[codec]} | {codec} | {d} | {r[d + '_GBps']:.1f} | {r['compression_ratio']:.3f} | {roof['achieved']:.1f} | "
                       f"{100 * roof['frac']:.1f} | {100 * roof['achieved'] / PEAK_NOMINAL:.1f} | {g.get(d + '_GBps', float('nan')):.1f} | "
                       f"{cpu.get(d + '_GBps', float('nan')):.2f} ({cpu.get('cores', '?')}) | {st.get(d + '_GBps', float('nan')):.2f} |")
    out.append("")
    out.append("Round trip (`value`) and end to end (`e2e`: pinned host buffers through `b200compStagerRoundTrip`), same run:\n")
    out.append("| codec | device round trip GB/s | e2e GB/s | e2e ms/step | reference GPU round trip | CPU arm round trip (`--impl reference`) | ratio vs reference encoder |")
    out.append("|---|---|---|---|---|---|---|")
    refc = codec_records(ref1) if ref1 else {}
    for codec, r in codec_records(b1).items():
        g = refgpu.get(codec, {})
        out.append(f"| {codec} | {r['value']:.1f} | {r['e2e']['value']:.1f} | {r['e2e']['ms_per_step']:.1f} | {g.get('value', float('nan')):.1f} | "
                   f"{refc.get(codec, {}).get('value', float('nan')):.2f} | {r['compression_ratio']:.3f} vs {g.get('compression_ratio', float('nan')):.3f} |")
    cc = b1["e2e"].get("copy_ceiling", {})
    if cc:
        out.append(f"\nPlain pinned copies on the same box: H2D {cc['h2d_GBps']:.1f}, D2H {cc['d2h_GBps']:.1f}, both at once "
                   f"{cc['duplex_GBps_each_way']:.1f} GB/s each way - the floor of an LZ4 e2e step is (raw + compressed bytes) / duplex rate.")
    # scaling
    out.append("\n### 5b. Scaling over the GPUs of one box\n")
    out.append("| N | LZ4 C2 round trip (weak) | e2e | Snappy C1 | Cascaded C3 | C4 LZ4 decompress-only 16 GiB (strong) | C4 size gather ms | copy ceiling (sum, each way) | CPU arm |")
    out.append("|---|---|---|---|---|---|---|---|---|")
    for n in (1, 2, 4, 8):
        b = load_line(f"r02z_bench_{n}gpu.json")
        c4 = load_line(f"r02z_c4_{n}gpu.json")
        rf = load_line(f"r02z_reference_{n}gpu.json")
        if not b:
            continue
        cd = b.get("codecs", {})
        ccn = b.get("e2e", {}).get("copy_ceiling", {})
        out.append(f"| {n} | {b['value']:.1f} | {b.get('e2e', {}).get('value', float('nan')):.1f} | {cd.get('snappy', {}).get('value', float('nan')):.1f} | "
                   f"{cd.get('cascaded', {}).get('value', float('nan')):.0f} | {c4['value'] if c4 else float('nan'):.0f} | "
                   f"{(c4 or {}).get('size_gather_ms') or 0:.1f} | {ccn.get('duplex_GBps_each_way', float('nan')):.0f} | "
                   f"{rf['value'] if rf else float('nan'):.2f} ({rf['cpu_baseline']['cores'] if rf else '?'} cores) |")
    out.append("\n(N = 1, 2, 8: the closing library.  N = 4, here and in 5c: records of the library before the last decoder and Cascaded changes "
               "(DESIGN §6 r2m-r2q) - Snappy decode 152 instead of 174 GB/s per GPU, Cascaded random-class compress 0.55 instead of 1.8 TB/s per GPU; "
               "LZ4 unchanged.)")
    # C5
    out.append("\n### 5c. C5 input-class sweep (16 GiB per class for the whole job; compress / decompress GB/s, ratio)\n")
    for n in (1, 2, 4, 8):
        rows = load_lines(f"r02z_c5_{n}gpu.jsonl")
        if not rows:
            continue
        out.append(f"N = {n}:\n")
        out.append("| class | LZ4 | Snappy | Cascaded (int32, R=1 D=1 bp) |" + (" CPU LZ4 | CPU Snappy | CPU Cascaded |" if n == 1 else ""))
        out.append("|---|---|---|---|" + ("---|---|---|" if n == 1 else ""))
        for k in ("zeros", "random", "text", "ints"):
            cells, cpus = [], []
            for codec in ("lz4", "snappy", "cascaded"):
                r = next((x for x in rows if x["input_class"] == k and x["codec"] == codec), None)
                if r is None:
                    cells.append("-")
                    cpus.append("-")
                    continue
                cells.append(f"{r['compress_GBps']:.0f} / {r['decompress_GBps']:.0f} ({r['compression_ratio']:.2f}; "
                             f"{100 * r['roofline_compress_frac']:.1f} % / {100 * r['roofline_decompress_frac']:.1f} %)")
                c = r.get("cpu")
                cpus.append(f"{c['compress_GBps']:.1f} / {c['decompress_GBps']:.1f}" if c else "-")
            out.append(f"| {k} | " + " | ".join(cells) + " |" + (" " + " | ".join(cpus) + " |" if n == 1 else ""))
        out.append("")
    out.append("(percentages: algorithmic bytes per second against the measured HBM peak, compress / decompress)\n")
    # cascaded variants
    rows = load_lines("r02z_casc_variants.jsonl")
    if rows:
        out.append("### 5d. Cascaded element types and option sets (1 GiB of the C3 column per type, N = 1)\n")
        out.append("| workload | compress GB/s | decompress GB/s | ratio | % of measured peak (c / d) |")
        out.append("|---|---|---|---|---|")
        for r in rows:
            out.append(f"| {r['config']['workload'].split(': ', 1)[-1]} | {r['compress_GBps']:.0f} | {r['decompress_GBps']:.0f} | "
                       f"{r['compression_ratio']:.2f} | {100 * r['roofline_compress']['frac']:.1f} / {100 * r['roofline_decompress']['frac']:.1f} |")
    text = "\n".join(out) + "\n"
    if "--write" in sys.argv:
        p = os.path.join(ROOT, "BASELINE.md")
        s = open(p).read()
        a = s.index("## 5. Results")
        s = s[:a] + "## 5. Results\n\n" + text
        open(p, "w").write(s)
    else:
        print(text)


if __name__ == "__main__":
    main()
