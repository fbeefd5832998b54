"""Oobleck VAE (oobleck_finetune.json widths) on the device: encode / decode time, algorithmic TFLOP/s, beside torch + cuDNN
running the oracle's functional restatement in bf16 autocast on the same GPU.  usage: python scripts/bench_vae.py [B] [T]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from ditsep_b200.models.oobleck import OobleckVAE
from oracle import oobleck_ref as O

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
T = int(sys.argv[2]) if len(sys.argv) > 2 else 32768
ch, cm, st, lat = 128, [1, 2, 4, 8, 16], [2, 4, 4, 8, 8], 128
sd = O.make_state_dict(ch, cm, st, lat, seed=7)
vae = OobleckVAE(ch, cm, st, lat, device="cuda").load_state_dict(sd)
sdc = {k: v.cuda() for k, v in sd.items()}


def flops(T, decoder):
    c = [ch * m for m in [1] + cm]
    f, t = 0.0, T
    for i, s in enumerate(st):
        f += 3 * 2.0 * t * c[i] * c[i] * 8          # three residual units: conv7 + conv1
        f += 2.0 * (t // s) * c[i] * c[i + 1] * 2 * s   # strided / transposed convolution
        t //= s
    return f


def timeit(fn, reps=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


x = 0.3 * torch.randn(B, 1, T, device="cuda")
z = torch.randn(2 * B, lat // 2, T // 2048, device="cuda")
t_enc = timeit(lambda: vae.encode(x, seed=1))
t_dec = timeit(lambda: vae.decode(z))
with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
    torch.backends.cudnn.benchmark = True
    t_enc_lib = timeit(lambda: O.encoder_forward(sdc, x, st))
    t_dec_lib = timeit(lambda: O.decoder_forward(sdc, z, st))
fe, fd = B * flops(T, False), 2 * B * flops(T, True)
print(f"Oobleck VAE, B={B} mixtures of {T} samples (decode: {2 * B} sources), bf16")
print(f"encode: {t_enc:8.2f} ms  {fe / t_enc / 1e9:7.1f} TFLOP/s | torch+cuDNN bf16 autocast {t_enc_lib:8.2f} ms")
print(f"decode: {t_dec:8.2f} ms  {fd / t_dec / 1e9:7.1f} TFLOP/s | torch+cuDNN bf16 autocast {t_dec_lib:8.2f} ms")
print(f"per mixture: encode {t_enc / B:.3f} ms + decode of 2 sources {t_dec / B:.3f} ms")
