"""Installs the UNMODIFIED reference files the CPU / stock-PyTorch baselines time into git-ignored baseline/_ref/
(BASELINE.md section 2): models/rendering.py, models/nerf.py, models/__init__.py and losses.py -- the only reference
modules that import with torch + einops alone (train.py / eval.py / datasets need pytorch_lightning, kornia, ... which
are not in the image).  The reference has no setup.py / pyproject.toml, so `pip install /root/reference` cannot work;
this file copy is the install.  baseline/_ref is git-ignored (no reference source enters the history) but NOT
gpurun-ignored, so it travels to the GPU box.  Called by __graft_entry__.build() when /root/reference is present."""
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
FILES = ("models/__init__.py", "models/rendering.py", "models/nerf.py", "losses.py", "LICENSE")


def install(src: str = "/root/reference") -> bool:
    if not os.path.isdir(src):
        return os.path.exists(os.path.join(DST, "models", "rendering.py"))
    for rel in FILES:
        d = os.path.join(DST, rel)
        os.makedirs(os.path.dirname(d), exist_ok=True)
        shutil.copyfile(os.path.join(src, rel), d)
    return True


if __name__ == "__main__":
    print("installed" if install() else "reference not available", DST)
