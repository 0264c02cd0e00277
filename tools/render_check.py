"""GPU check of the device path tracer + progressive training loop on the Cornell box:
(1) BSDF-only passes, (2) the SDMM loop; both must converge to the same image (unbiased MIS),
the guided one with less variance per sample.  Writes gpurun_out/render_*.npy/png."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from sdmm_mitsuba_b200 import scenes
from sdmm_mitsuba_b200.integrator import SDMMIntegrator

def save_png(path, img):
    try:
        from PIL import Image
        x = np.clip(img / (1 + img), 0, 1) ** (1 / 2.2)
        Image.fromarray((x.transpose(1, 2, 0) * 255).astype(np.uint8)).save(path)
    except Exception as e:
        print("png skipped:", e)

W, H = int(os.environ.get("W", 320)), int(os.environ.get("H", 180))
K = int(os.environ.get("K", 16)); MD = int(os.environ.get("MAXDEPTH", 5))
SPI = int(os.environ.get("SPI", 4)); SC = int(os.environ.get("SAMPLES", 64))
quads, cam = scenes.cornell_box(W, H)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
# reference: many BSDF-only passes
ref = SDMMIntegrator(quads, cam, K=K, max_depth=MD, samples_per_iteration=SPI, seed=99)
t0 = time.time()
REF_SPP = int(os.environ.get("REF_SPP", 1024))
for i in range(REF_SPP // 64):
    ref.iteration = 1000 + i
    ref.render_pass(64, collect=False, guided=False)
    ref.total_spp += 64
torch.cuda.synchronize()
print("bsdf-only reference %d spp: %.2f s" % (REF_SPP, time.time() - t0))
img_ref = ref.image()
# equal-sample BSDF-only image
b = SDMMIntegrator(quads, cam, K=K, max_depth=MD, samples_per_iteration=SPI, seed=5)
for i in range(SC // SPI):
    b.iteration = i
    b.render_pass(SPI, collect=False, guided=False); b.total_spp += SPI
img_b = b.image()
g = SDMMIntegrator(quads, cam, K=K, max_depth=MD, samples_per_iteration=SPI, seed=5)
t0 = time.time()
img_g = g.render(SC, log=lambda d: print(json.dumps(d)))
torch.cuda.synchronize()
print("guided loop: %.2f s" % (time.time() - t0))
def mse(a): return float(((a - img_ref) ** 2).mean()), float(np.abs(a.mean() - img_ref.mean()) / img_ref.mean())
print("mean radiance ref/bsdf/guided:", img_ref.mean(), img_b.mean(), img_g.mean())
print("MSE vs ref (and rel. mean error): bsdf", mse(img_b), "guided", mse(img_g))
# last (fully guided, non-training) iterations only
for name, im in (("ref", img_ref), ("bsdf", img_b), ("guided", img_g)):
    np.save(os.path.join(ROOT, "gpurun_out", "render_%s.npy" % name), im.astype(np.float16))
    save_png(os.path.join(ROOT, "gpurun_out", "render_%s.png" % name), im)
