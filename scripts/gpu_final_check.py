"""Last check of the committed tree on one B200, one Python process (a fresh box pays ~1 min per `import torch`):
smoke() -> the data-loader (gather) path against the resident path -> an out-of-range index must come back as an
error, not a crash -> a short bench.  Every stage appends to gpurun_out/final_check/log.txt as it finishes."""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "gpurun_out", "final_check")
os.makedirs(OUT, exist_ok=True)
T0 = time.time()


def log(msg: str) -> None:
    with open(os.path.join(OUT, "log.txt"), "a") as f:
        f.write(f"[{time.time() - T0:6.1f}s] {msg}\n")
    print(msg, flush=True)


def main() -> None:
    import torch
    log(f"torch {torch.__version__} imported, device {torch.cuda.get_device_name(0)}")
    import __graft_entry__ as g
    g.smoke()
    log("smoke ok")
    from pytorch_mnist_ddp_b200 import _native as N
    from pytorch_mnist_ddp_b200.engine import B200Backend
    from pytorch_mnist_ddp_b200.parallel.dist import DistInfo
    from pytorch_mnist_ddp_b200.utils.data import synthetic_split
    train, test = synthetic_split(True, 4096), synthetic_split(False, 256)
    idx = torch.randperm(4096, generator=torch.Generator().manual_seed(3)).to(torch.int32)
    losses = {}
    for path in ("resident", "gather"):
        torch.manual_seed(1)
        be = B200Backend(train, test, DistInfo(False, 0, 1, 0, ""), lr=1.0, seed=1, max_batch=256)
        be.begin_epoch(idx)
        steps = []
        for k in range(8):
            if path == "resident":
                steps.append(be.step(200))
            else:
                steps.append(be.step_gather(train.images, train.labels, idx[k * 200:(k + 1) * 200].contiguous()))
        be.finish()
        losses[path] = [be.engine.read_loss(s) for s in steps]
        if path == "gather":
            bad = idx[:200].clone()
            bad[17] = 4096
            try:
                be.step_gather(train.images, train.labels, bad)
                log("FAIL: an out-of-range index was accepted")
            except N.NativeError as e:
                log(f"out-of-range index rejected: {e}")
        be.engine.close()
    same = losses["resident"] == losses["gather"]
    log(f"gather path == resident path over 8 steps: {same}  {losses['gather'][:3]}")
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "100", "--warmup", "5", "--epochs", "1"],
                         capture_output=True, text=True, cwd=ROOT)
    open(os.path.join(OUT, "bench.json"), "w").write(res.stdout)
    open(os.path.join(OUT, "bench.err"), "w").write(res.stderr[-4000:])
    try:
        d = json.loads(res.stdout.strip().splitlines()[-1])
        log(f"bench rc {res.returncode}: {d['ms_per_step'] * 1000:.1f} us/step, e2e {d['e2e']['value']:.0f} img/s, "
            f"golden {d['checks']['golden_losses'].get('status')} ok {d['checks']['ok']}")
    except Exception as e:  # noqa: BLE001
        log(f"bench rc {res.returncode}: unparsable output ({e})")


if __name__ == "__main__":
    try:
        main()
    except Exception as e:  # noqa: BLE001
        import traceback
        log("EXCEPTION " + "".join(traceback.format_exception(e))[-3000:])
        raise
