"""Stage the UNMODIFIED reference modules under ``oracle/_ref/`` (test / baseline infrastructure only).

    python oracle/stage_ref.py            # needs /root/reference (the build container)

``/root/reference`` does not exist on the GPU box, but the tree -- including the git-ignored ``oracle/_ref/`` -- is what
``gpurun`` and the driver ship there.  ``__graft_entry__.build()`` runs this recipe wherever the reference is readable, so
the GPU-side tests and ``bench.py``'s reference legs run the reference's own code (``model/tacotron2.py`` and the files it
imports, all torch-only; plus ``model/hifi_gan.py`` + ``model/resblock.py`` for the vocoder row) instead of a restatement.
The files are copied byte for byte; nothing under ``oracle/_ref/`` is committed (``.gitignore``) and nothing in the product
(``tacotron2_b200/``) may import it.  ``MANIFEST.json`` records the sha256 of every staged file.
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
SRC = os.environ.get("T2_REFERENCE_ROOT", "/root/reference")
FILES = [
    "model/__init__.py", "model/tacotron2.py", "model/encoder.py", "model/decoder.py", "model/attention.py", "model/postnet.py",
    "model/modules.py", "model/embedding_encoder.py", "model/hifi_gan.py", "model/resblock.py", "datasets/tts_dataloader.py",
]
CONFIG_DIR = "config"   # the reference's JSON configs (tests map every one of them through TTSModel.from_config)


def stage(verbose: bool = True) -> bool:
    """Returns True when oracle/_ref is complete afterwards (staged now or already there)."""
    if not os.path.isdir(SRC):
        ok = os.path.exists(os.path.join(DST, "MANIFEST.json"))
        if verbose:
            print(f"stage_ref: {SRC} absent; oracle/_ref {'present' if ok else 'MISSING'}")
        return ok
    manifest = {}
    for rel in FILES:
        s, d = os.path.join(SRC, rel), os.path.join(DST, rel)
        os.makedirs(os.path.dirname(d), exist_ok=True)
        shutil.copyfile(s, d)
        manifest[rel] = hashlib.sha256(open(d, "rb").read()).hexdigest()
    cdir = os.path.join(SRC, CONFIG_DIR)
    os.makedirs(os.path.join(DST, CONFIG_DIR), exist_ok=True)
    for name in sorted(os.listdir(cdir)):
        if name.endswith(".json"):
            shutil.copyfile(os.path.join(cdir, name), os.path.join(DST, CONFIG_DIR, name))
            manifest[f"{CONFIG_DIR}/{name}"] = hashlib.sha256(open(os.path.join(DST, CONFIG_DIR, name), "rb").read()).hexdigest()
    open(os.path.join(DST, "datasets", "__init__.py"), "a").close()
    json.dump(manifest, open(os.path.join(DST, "MANIFEST.json"), "w"), indent=1, sort_keys=True)
    if verbose:
        print(f"stage_ref: {len(manifest)} files -> {DST}")
    return True


if __name__ == "__main__":
    sys.exit(0 if stage() else 1)
