"""CPU checks of bench.py's bookkeeping (no GPU work): per-step percentiles, the stage byte / flop models cover every
stage the library times, and `roofline.traffic` is only quoted from an ncu capture of the kernel source as it is now."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def test_percentiles():
    p = bench.percentiles([float(i) for i in range(1, 102)])
    assert p == {"median": 51.0, "p10": 11.0, "p90": 91.0, "n": 101}
    assert bench.percentiles([]) is None


def test_stage_model_names_match_the_library():
    from gaussian_splatting_b200 import _lib, build
    build.build()
    lib = _lib.load()
    names = {lib.gsr_stage_name(i).decode() for i in range(_lib.STAGE_COUNT)}
    work = dict(P=1000, V=800, R=6000, T=300, N=76800, Wf=10, Wb=10, Bp=5, chunks=15, world=1)
    single = bench.stage_model(work)
    assert set(single) == names - {"exchange"}
    work["world"] = 8
    multi = bench.stage_model(work)
    assert set(multi) == names and multi["exchange"][1] == (16 * 8 + 48 + 12 + 236) * 1000
    assert multi["preprocess_bwd"][1] < single["preprocess_bwd"][1]      # the sender branch publishes 64 B records
    assert single["depth_sort"] == ("hbm", 72 * 1000)


def test_ncu_evidence_is_tied_to_the_current_kernel_source(tmp_path, monkeypatch):
    sha = bench.source_sha1("render_bwd.cu")
    entries = [dict(kernel="render_bwd_kernel<1>", dram_read_MB=130.0, dram_write_MB=10.0, issue_active_pct=88.0, src_sha1=sha),
               dict(kernel="render_fwd_kernel", dram_read_MB=90.0, dram_write_MB=27.0, issue_active_pct=87.0, src_sha1="stale")]
    prof = tmp_path / "profiles" / "r99"
    prof.mkdir(parents=True)
    (prof / "r99_kernels.json").write_text(json.dumps(entries))
    csrc = tmp_path / "gaussian_splatting_b200" / "csrc"
    csrc.mkdir(parents=True)
    for f in ("render_bwd.cu", "render_fwd.cu"):
        (csrc / f).write_bytes(open(os.path.join(ROOT, "gaussian_splatting_b200", "csrc", f), "rb").read())
    monkeypatch.setattr(bench, "ROOT", str(tmp_path))
    ev = bench.ncu_evidence("render_bwd")
    assert ev["traffic"] == 140_000_000 and ev["src_sha1"] == sha and ev["source"].endswith("r99_kernels.json")
    assert bench.ncu_evidence("render_fwd") is None          # captured from another version of the kernel
    assert bench.ncu_evidence("tile_fill") is None
