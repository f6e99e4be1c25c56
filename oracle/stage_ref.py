"""Stages the REFERENCE's own modules for this path under oracle/_ref/ (TEST INFRASTRUCTURE).

    python oracle/stage_ref.py            # run in the build container, where /root/reference exists

The reference (kan-bayashi/espnet v0.9.5) is pure Python: there is nothing to compile, so "building"
oracle/_ref means copying, unmodified and under their own package paths, the few modules that hold the
reference's side of the boundary:

    espnet/nets/scorer_interface.py, beam_search.py, batch_beam_search.py, e2e_asr_common.py (end_detect)
    espnet/nets/ctc_prefix_score.py, scorers/ctc.py, scorers/length_bonus.py
    espnet/nets/pytorch_backend/ctc.py (+ nets_utils.py for to_device), espnet2/asr/ctc.py

oracle/_ref/ is git-ignored (the sources never enter this repository's history) but not gpurun-ignored: like
the built .so files it travels to the GPU box with the working tree, where /root/reference does not exist.
Only tests/, __graft_entry__.smoke() and bench.py's reference legs import it, as the checker and the timed
baseline; the product package never does (tests/test_host_logic.py::test_product_never_imports_the_oracle).
`editdistance` (imported by e2e_asr_common, never called by end_detect) is the two-line stub the golden
generators already use.
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
FILES = [
    "espnet/nets/scorer_interface.py",
    "espnet/nets/beam_search.py",
    "espnet/nets/batch_beam_search.py",
    "espnet/nets/e2e_asr_common.py",
    "espnet/nets/ctc_prefix_score.py",
    "espnet/nets/scorers/ctc.py",
    "espnet/nets/scorers/length_bonus.py",
    "espnet/nets/pytorch_backend/ctc.py",
    "espnet/nets/pytorch_backend/nets_utils.py",
    "espnet/nets/pytorch_backend/rnn/decoders.py",      # recognize_beam_batch: the multi-utterance search loop (8f4)
    "espnet/nets/pytorch_backend/rnn/attentions.py",
    "espnet2/asr/ctc.py",
]
PACKAGES = ["espnet", "espnet/nets", "espnet/nets/scorers", "espnet/nets/pytorch_backend", "espnet/nets/pytorch_backend/rnn",
            "espnet2", "espnet2/asr"]
# e2e_asr_common imports the transducer search at module level (never used by end_detect); its dependency tree
# (transducer utils, espnet2 decoders) is not part of this path, so a one-line stand-in is written instead
STUBS = {"espnet/nets/beam_search_transducer.py": "class BeamSearchTransducer:  # stand-in, see oracle/stage_ref.py\n    pass\n"}


def stage(ref_root="/root/reference", force=False):
    """Returns DST when the staged tree is complete (copying it first if the reference is present)."""
    have = all(os.path.exists(os.path.join(DST, f)) for f in FILES)
    if not os.path.isdir(ref_root):
        return DST if have else None
    if have and not force:
        return DST
    for pkg in PACKAGES:
        os.makedirs(os.path.join(DST, pkg), exist_ok=True)
        init = os.path.join(DST, pkg, "__init__.py")
        if not os.path.exists(init):
            open(init, "w").close()  # empty package markers (the reference's own __init__ files are empty or absent)
    for f in FILES:
        shutil.copyfile(os.path.join(ref_root, f), os.path.join(DST, f))
    for f, text in STUBS.items():
        with open(os.path.join(DST, f), "w") as fh:
            fh.write(text)
    shutil.copyfile(os.path.join(HERE, "..", "tests", "golden", "stubs", "editdistance.py"),
                    os.path.join(DST, "editdistance.py"))
    return DST


def path():
    """The staged tree, or None (tests skip, bench legs fall back to the port)."""
    return DST if all(os.path.exists(os.path.join(DST, f)) for f in FILES) else None


def activate():
    """Puts the staged reference on sys.path (in front).  Returns True if it is there."""
    p = path()
    if p is None:
        return False
    if p not in sys.path:
        sys.path.insert(0, p)
    return True


if __name__ == "__main__":
    print(stage(force="--force" in sys.argv))
