#!/usr/bin/env python
"""Per-kernel roofline table of one training step, regenerated from the committed evidence.

    scripts/roofline.py [--batch 200] [--launches profiles/r2/gpus1/step_launches_warm.csv]
                        [--raw profiles/r2/gpus1/step_full_raw.csv] > profiles/r2/roofline_table.md

For every kernel of the step: useful FLOPs and the bytes of every tensor it has to touch once (derived here from the
model's dimensions, mnist_ddp.py:42-47, and the formats the engine keeps them in), the roofline time
max(FLOPs / TF32 peak, bytes / HBM copy bandwidth) with the peaks of MEASURED_PEAKS.json (TF32 dense = half of the
measured bf16 rate), the warm duration from the per-launch ncu log, and -- where the `--set full` raw page has the
kernel -- the DRAM bytes actually moved, tensor-pipe and issue utilisation.
"""
import argparse
import csv
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def peaks():
    hbm, bf16 = 6545.0, 1662.0  # GB/s, TFLOP/s: the values recorded during the round, used when the driver's file is absent
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        hbm = float(p.get("hbm_gbs", hbm))
        bf16 = float(p.get("bf16_tflops", bf16))
    except Exception:
        pass
    return hbm, bf16 / 2.0


def model(B: int, nsm: int = 148, splits: int = 36):
    """name -> (pattern in the ncu kernel name, useful FLOPs, [(tensor, bytes)])"""
    img, a1, pooled, amax = 784 * B, 26 * 26 * 32 * 4 * B, 9216 * 4 * B, 9216 * B
    w1, w2, wf1, wf2 = 32 * 9 * 4 + 128, 64 * 288 * 4 + 256, 128 * 9216 * 4, 10 * 128 * 4
    h = 128 * 4 * B
    c2part = nsm * (18432 + 64) * 4
    c1part = 2 * B * 320 * 4
    fc_bucket = (128 * 9216 + 128 + 10 * 128 + 10)
    conv_bucket = 19456
    f_conv1, f_conv2 = 2 * 676 * 32 * 9 * B, 2 * 576 * 64 * 288 * B
    f_dgrad = 2 * 676 * 32 * 576 * B
    f_fc1 = 2 * 9216 * 128 * B
    return [
        ("conv1+conv2 fwd", "conv_fwd_tc_kernel", f_conv1 + f_conv2,
         [("images u8", img), ("weights", w1 + w2), ("a1 out", a1), ("pooled out", pooled), ("arg-max out", amax)]),
        ("fc1 fwd (split-K)", "tc_gemm_kernel<64, 0, 0", f_fc1,
         [("pooled", pooled), ("fc1.weight", wf1), ("partials out", splits * h)]),
        ("head_fused", "head_fused_kernel", 2 * 128 * 10 * 2 * B,
         [("partials", splits * h), ("h, dh, dlogits out", 2 * h + 64 * B), ("fc2.weight", wf2)]),
        ("fc1 dgrad", "tc_gemm_kernel<128, 0, 1", f_fc1,
         [("dh", h), ("fc1.weight", wf1), ("pooled (mask)", pooled), ("dp out", pooled)]),
        ("head_wgrad", "head_wgrad_kernel", 2 * 128 * 10 * B, [("h, dh, dlogits", 2 * h + 64 * B)]),
        ("fc1 wgrad", "tc_gemm_kernel<64, 1, 1", f_fc1, [("dh", h), ("pooled", pooled), ("dW out", wf1)]),
        ("adadelta fc bucket", "adadelta_kernel", 10 * fc_bucket, [("g, v, u, p in; v, u, p out", 28 * fc_bucket)]),
        ("conv2 wgrad", "conv2_wgrad_tc_kernel", f_conv2,
         [("dp", pooled), ("arg-max", amax), ("a1", a1), ("per-CTA partial rows out", c2part)]),
        ("conv2 dgrad", "conv2_dgrad_tc_kernel", f_dgrad,
         [("dp", pooled), ("arg-max", amax), ("conv2.weight", w2), ("a1 (ReLU mask)", a1), ("da1 out", a1)]),
        ("conv2 partial reduce", "reduce_segs_kernel", c2part // 4, [("partial rows", c2part), ("gradient out", 18688 * 4)]),
        ("conv1 wgrad", "conv1_wgrad_kernel", 2 * 676 * 32 * 9 * B,
         [("da1", a1), ("images u8", img), ("partial rows out", c1part)]),
        ("adadelta conv bucket (end of step)", "adadelta_finalize_kernel", 10 * conv_bucket,
         [("g, v, u, p", 28 * conv_bucket), ("conv1 group rows", 13 * 320 * 4)]),
    ]


def read_launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    lines = [lines[0]] + [l for l in lines[1:] if not l.startswith('"ID"')]
    rows = [(r["Kernel Name"], float(r["Metric Value"].replace(",", "")) / 1000.0) for r in csv.DictReader(lines)
            if r.get("Metric Name") == "gpu__time_duration.sum"]
    starts = [i for i, (n, _) in enumerate(rows) if "conv_fwd_tc_kernel" in n]
    return rows[starts[-2]:starts[-1]] if len(starts) >= 2 else rows


def read_raw(path):
    rows = list(csv.reader(open(path)))
    hdr = rows[0]
    col = {c: i for i, c in enumerate(hdr)}
    want = {"dram_rd": "dram__bytes_read.sum", "dram_wr": "dram__bytes_write.sum",
            "tensor": "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
            "issue": "smsp__issue_active.avg.pct_of_peak_sustained_active", "regs": "launch__registers_per_thread",
            "smem": "launch__shared_mem_per_block_dynamic"}
    units = rows[1]
    out = {}
    for r in rows[2:]:
        rec = {}
        for k, c in want.items():
            if c in col and r[col[c]] != "":
                v = float(r[col[c]].replace(",", ""))
                u = units[col[c]]
                if k.startswith("dram"):
                    v *= {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0}.get(u, 1.0)
                rec[k] = v
        out[r[col["Kernel Name"]]] = rec
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=200)
    ap.add_argument("--launches", default=os.path.join(ROOT, "profiles/r2/gpus1/step_launches_warm.csv"))
    ap.add_argument("--raw", default=os.path.join(ROOT, "profiles/r2/gpus1/step_full_raw.csv"))
    a = ap.parse_args()
    hbm, tf32 = peaks()
    launches = read_launches(a.launches)
    raw = read_raw(a.raw) if os.path.isfile(a.raw) else {}
    print(f"Roofline of one training step, batch {a.batch}, one B200.  Peaks: HBM copy {hbm:.0f} GB/s, TF32 dense "
          f"{tf32:.0f} TFLOP/s (half of the measured bf16 rate).  `bytes` = every tensor the kernel must touch once, in "
          f"the formats the engine keeps; `DRAM moved` = ncu, cold caches (tensors a previous kernel left in the 126 MB "
          f"L2 do not show up there).\n")
    print("| kernel | useful GFLOP | bytes MB | roofline µs (bound) | warm µs | fraction | DRAM moved MB (ncu) | tensor pipe % of active cycles | issue slots % of active cycles |")
    print("|---|---|---|---|---|---|---|---|---|")
    tot_roof = tot_warm = 0.0
    for name, pat, flops, tensors in model(a.batch):
        by = sum(b for _, b in tensors)
        t_f, t_b = flops / (tf32 * 1e12) * 1e6, by / (hbm * 1e9) * 1e6
        roof, bound = max(t_f, t_b), ("tensor" if t_f > t_b else "memory")
        warm = next((us for n, us in launches if pat in n), None)
        rec = next((v for n, v in raw.items() if pat in n), {})
        dram = (rec.get("dram_rd", 0) + rec.get("dram_wr", 0)) / 1e6 if rec else None
        tot_roof += roof
        tot_warm += warm or 0.0
        f = lambda v, p=1: "–" if v is None else f"{v:.{p}f}"
        print(f"| {name} | {flops / 1e9:.2f} | {by / 1e6:.1f} | {roof:.1f} ({bound}) | {f(warm)} | "
              f"{'–' if not warm else f'{100 * roof / warm:.0f} %'} | {f(dram)} | {f(rec.get('tensor'), 0) if rec else '–'} | "
              f"{f(rec.get('issue'), 0) if rec else '–'} |")
    print(f"| **sum** | | | **{tot_roof:.1f}** | **{tot_warm:.1f}** | {100 * tot_roof / tot_warm:.0f} % | | | |")
    print("\nTensors per kernel:\n")
    for name, _, _, tensors in model(a.batch):
        print(f"* {name}: " + ", ".join(f"{t} {b / 1e6:.2f} MB" for t, b in tensors))


if __name__ == "__main__":
    main()
