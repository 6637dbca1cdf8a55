"""Bring-up timing of the config-5 training step on one GPU: ours (dx_train_step + dx_adam_step), per-kernel profile, and
the unmodified reference classes run eagerly on the same GPU (forward + loss + backward + Adam)."""
import ctypes
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from daft_exprt_b200 import _lib, synthetic  # noqa: E402
from daft_exprt_b200.hparams import default_hparams  # noqa: E402
from daft_exprt_b200.weights import make_state_dict  # noqa: E402
from daft_exprt_b200.train import TrainEngine, loss_dict, flops_per_step  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    hp = default_hparams()
    sd = make_state_dict(hp, 1234)
    batch = synthetic.forward_batch(hp, batch=B, lmin=20, lmax=150, seed=42)
    dev = tuple(t.cuda() for t in batch)
    frames = int(batch[9].sum())
    mode = sys.argv[2] if len(sys.argv) > 2 else "fp32"
    eng = TrainEngine(hp, sd, gemm=mode)
    print("GEMM mode:", mode)
    for kv in os.environ.get("DX_KNOBS", "").split(","):        # e.g. DX_KNOBS=18=148
        if kv:
            k, v = kv.split("=")
            _lib.load().dx_set_debug(int(k), int(v))
            print("knob", k, "=", v)
    for it in range(1, 4):
        eng.step(dev, iteration=it)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 10
    e0.record()
    for it in range(4, 4 + n):
        eng.step(dev, iteration=it)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    fl = flops_per_step(batch[5].tolist(), batch[9].tolist())
    print(f"ours: B={B} frames={frames} step {ms:.3f} ms  {frames / ms * 1e3 / 1e6:.3f} M frames/s  {fl / ms / 1e9:.1f} TFLOP/s (fwd+bwd algorithmic)")
    print("losses", loss_dict(eng.losses))
    lib = _lib.load()
    lib.dx_profile_enable(1)
    eng.step(dev, iteration=100)
    buf = ctypes.create_string_buffer(1 << 16)
    lib.dx_profile_report(buf, len(buf))
    lib.dx_profile_enable(0)
    print(buf.value.decode())
    print("workspace MB", eng._ws.numel() / 1e6)
    if os.environ.get("DX_NO_REF"):
        return
    # the reference, eager, on this GPU
    try:
        from oracle import ref_harness as R
        import copy
        ref = R.build_reference_model(hp, sd).cuda()
        import daft_exprt.loss as ref_loss
        crit = ref_loss.DaftExprtLoss("cuda", copy.deepcopy(hp))
        opt = torch.optim.Adam(ref.parameters(), betas=tuple(hp.betas), eps=hp.epsilon, weight_decay=hp.weight_decay)
        torch.backends.cuda.matmul.allow_tf32 = False
        torch.backends.cudnn.allow_tf32 = False

        def ref_step(it):
            inputs = tuple(t.clone() for t in dev)
            targets = (inputs[1], inputs[3], inputs[4], inputs[8], inputs[9], inputs[10])
            opt.zero_grad(set_to_none=True)
            out = ref(inputs)
            loss, _ = crit(out, targets, it)
            loss.backward()
            opt.step()
            return loss
        for it in range(1, 3):
            ref_step(it)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for it in range(3, 6):
            ref_step(it)
        torch.cuda.synchronize()
        rms = (time.perf_counter() - t0) / 3 * 1e3
        print(f"reference eager fp32 on this GPU (eval mode, TF32 off): step {rms:.1f} ms  -> ours is {rms / ms:.1f} x faster")
    except Exception as e:      # noqa: BLE001
        print("reference eager GPU step failed:", type(e).__name__, e)


if __name__ == "__main__":
    main()
