"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/rt_api.h
declares, refuses to run without a GPU, and its host-only pieces (PPM writer, tile queue) work."""
import ctypes
import multiprocessing as mp
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "rt_api.h")).read()
    return sorted(set(re.findall(r"RT_EXPORT\s+[\w\s\*]+?\b(rt_\w+)\s*\(", text)))


def test_header_symbols_are_exported():
    from raytracinginoneweekend_b200 import backend
    lib = backend.load_library()
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/rt_api.h but not exported"
    assert sorted(backend.EXPORTS) == names, "backend.EXPORTS must list exactly the header's entry points"


def test_python_flags_match_the_header():
    """backend.FLAG_* / multigpu.FLAG_OVERLAP are the RT_FLAG_* values of include/rt_api.h."""
    from raytracinginoneweekend_b200 import backend, multigpu
    text = open(os.path.join(ROOT, "include", "rt_api.h")).read()
    header = {m.group(1): 1 << int(m.group(2)) for m in re.finditer(r"\bRT_FLAG_(\w+)\s*=\s*1u\s*<<\s*(\d+)", text)}
    assert len(header) >= 7 and len(set(header.values())) == len(header)
    mine = {"EXACT": backend.FLAG_EXACT, "ACCEL_BRUTE": backend.FLAG_BRUTE, "ACCEL_LBVH": backend.FLAG_LBVH,
            "FILTER_FMA": backend.FLAG_FILTER_FMA, "FILTER_CHECK": backend.FLAG_FILTER_CHECK,
            "BOOK_OPTICS": backend.FLAG_BOOK_OPTICS, "OVERLAP": backend.FLAG_OVERLAP,
            "TEST_STRAGGLERS": backend.FLAG_TEST_STRAGGLERS}
    assert mine == {k: header[k] for k in mine}
    assert multigpu.FLAG_OVERLAP == header["OVERLAP"]


def test_no_cpu_fallback():
    """Without a usable GPU rt_create must fail loudly (RT_ERR_NO_DEVICE), never render on the CPU."""
    import torch
    from raytracinginoneweekend_b200 import backend
    if torch.cuda.is_available():
        pytest.skip("a GPU is present; the refusal path is covered on the CPU runner")
    with pytest.raises(backend.RtError, match="no usable CUDA device|no CPU fallback"):
        backend.Backend(0)


def test_static_entry_point_refuses_without_a_gpu():
    """rt_closest_hits_static (no context, no allocation): without a usable GPU it fails with a message, before any launch;
    its size limits are checked before the device is touched."""
    import numpy as np
    import torch
    from raytracinginoneweekend_b200 import backend
    sph, ray = np.float32([[0, 0, -1, .5]]), np.float32([[0, 0, 0, 0, 0, -1]])
    with pytest.raises(backend.RtError, match="RT_STATIC_MAX_SPHERES"):
        backend.closest_hits_static(np.tile(sph, (backend.STATIC_MAX_SPHERES + 1, 1)), ray)
    with pytest.raises(backend.RtError, match="RT_STATIC_MAX_RAYS"):
        backend.closest_hits_static(sph, np.tile(ray, (backend.STATIC_MAX_RAYS + 1, 1)))
    if torch.cuda.is_available():
        pytest.skip("a GPU is present; the refusal path is covered on the CPU runner")
    with pytest.raises(backend.RtError, match="rt_closest_hits_static: cuda"):
        backend.closest_hits_static(sph, ray)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "raytracinginoneweekend_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".inc", ".hxx", ".cpp", ".cxx")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f"{f} imports the oracle"
                assert not re.search(r"#\s*include\s*[\"<][^\">]*oracle", src), f"{f} includes oracle code"
                assert "librt_oracle" not in src and "libref_oracle" not in src, f"{f} loads an oracle library"


def test_write_ppm_matches_reference_writer(tmp_path, ref):
    """rt_write_ppm vs app::save_to_file (src/main.cxx:87-101): identical bytes."""
    from raytracinginoneweekend_b200 import backend
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, (19, 33, 3), dtype=np.uint8)
    a, b = str(tmp_path / "a.ppm"), str(tmp_path / "b.ppm")
    backend.write_ppm(a, img)
    assert ref.save_ppm(b, img) == 0
    assert open(a, "rb").read() == open(b, "rb").read()


def test_write_ppm_layout(tmp_path):
    from raytracinginoneweekend_b200 import backend
    img = np.arange(2 * 3 * 3, dtype=np.uint8).reshape(2, 3, 3)
    p = str(tmp_path / "x.ppm")
    backend.write_ppm(p, img)
    assert open(p, "rb").read() == b"P6\n3 2\n255\n" + img.tobytes()
    with pytest.raises(backend.RtError, match="bad file"):
        backend.write_ppm(str(tmp_path / "nope" / "x.ppm"), img)


def _puller(name, slot, n, out):
    from raytracinginoneweekend_b200.backend import TileQueue
    q = TileQueue(name, create=False)
    got = [q.next(slot, 1) for _ in range(n)]
    q.close()
    out.put(got)


def test_tile_queue_is_shared_and_atomic_across_processes():
    from raytracinginoneweekend_b200.backend import TileQueue
    name = f"/rtiow_test_{os.getpid()}"
    q = TileQueue(name, create=True)
    try:
        q.reset(3, 100)
        assert q.next(3, 0) == 100 and q.next(0, 0) == 0
        ctx = mp.get_context("spawn")
        out = ctx.Queue()
        procs = [ctx.Process(target=_puller, args=(name, 3, 500, out)) for _ in range(3)]
        for p in procs:
            p.start()
        got = sum((out.get(timeout=120) for _ in procs), [])
        for p in procs:
            p.join(timeout=60)
        assert sorted(got) == list(range(100, 1600))      # every value handed out exactly once
        assert q.next(3, 0) == 1600
        with pytest.raises(Exception):
            q.next(64, 1)                                  # slot out of range
    finally:
        q.close()
