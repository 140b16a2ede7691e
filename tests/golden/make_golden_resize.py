"""Generate `golden_resize.pt` from the UNMODIFIED reference's DynamicResize (acai_omr/utils/utils.py:334-370), preceded by
v2.ToImage + v2.ToDtype(float32, scale=True) for the uint8 cases (omr_teacher_force_train.py:278-282).

Run in the build container only (needs /root/reference and torchvision):  python tests/golden/make_golden_resize.py
`acai_omr.utils.utils` imports matplotlib and `acai_omr/__init__.py` imports flask, neither installed: both are stubbed
(empty modules); nothing of them is used by DynamicResize."""
import os
import sys
import tempfile

import torch

REF = "/root/reference"
stub = tempfile.mkdtemp()
for pkg, body in (("flask", "class Flask: pass\n"), ("matplotlib", "")):
    os.makedirs(os.path.join(stub, pkg))
    with open(os.path.join(stub, pkg, "__init__.py"), "w") as f:
        f.write(body)
open(os.path.join(stub, "matplotlib", "pyplot.py"), "w").close()
sys.path[:0] = [stub, REF]
sys.dont_write_bytecode = True
os.chdir(REF)

from torchvision.transforms import v2  # noqa: E402

from acai_omr.utils.utils import DynamicResize  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_resize.pt")

# (H, W, dtype, patch_size, max_seq_len, pe_max_height, pe_max_width, crop_imgs): down- and up-scaling on both axes, portrait and
# landscape, non-integer ratios, the centre crop on either axis (odd offsets), small token budgets to keep the file small
CASES = [
    (90, 700, torch.uint8, 16, 64, 60, 200, False),      # -> 32 x 224, downscale ~2.8x / 3.1x
    (37, 53, torch.uint8, 16, 64, 60, 200, False),       # -> 128 x 128, upscale
    (300, 120, torch.float32, 16, 64, 60, 200, False),   # portrait -> 128 x 64
    (61, 333, torch.float32, 16, 96, 60, 200, False),    # -> 64 x 320: up in height, about 1:1 in width
    (100, 950, torch.uint8, 16, 100, 60, 200, True),     # -> 48 x 432, no crop needed
    (40, 500, torch.uint8, 8, 120, 60, 20, True),        # -> 24 x 288, width cropped to 160 (offset 64)
    (500, 70, torch.float32, 8, 150, 9, 200, True),      # portrait -> 224 x 32, height cropped to 72 (offset 76)
    (333, 333, torch.uint8, 16, 50, 60, 200, False),     # square -> 112 x 112
    (17, 2000, torch.uint8, 16, 1024, 60, 100, True),    # aspect 117 -> 32 x 3744 cropped to 1600 wide
]

entries = []
g = torch.Generator().manual_seed(7)
for H, W, dt, P, N, ph, pw, crop in CASES:
    if dt == torch.uint8:
        img = torch.randint(0, 256, (1, H, W), generator=g, dtype=torch.uint8)
        img[:, H // 3: H // 3 + 2, :] = 0                     # a staff line
        x = v2.ToDtype(torch.float32, scale=True)(v2.ToImage()(img))
    else:
        img = torch.rand(1, H, W, generator=g)
        x = img
    out = DynamicResize(P, N, ph, pw, crop)(x)
    entries.append(dict(img=img, args=(P, N, ph, pw, crop), out=torch.as_tensor(out).clone()))
    print(tuple(img.shape), img.dtype, "->", tuple(out.shape))
torch.save(entries, OUT)
print("wrote", OUT, os.path.getsize(OUT), "bytes")
