"""DecodeJpeg (EX/common/GraphBuilder.java:41-47) without a GPU: the oracle against the libjpeg-turbo fixtures, the
kernels' per-thread functions stepped on the CPU against the same fixtures, and the host-side header reader of the
C ABI (ftf_jpeg_info) with its error behaviour."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def cases():
    z = np.load(os.path.join(ROOT, "tests", "golden", "jpeg", "cases.npz"))
    names = sorted(k[:-4] for k in z.files if k.endswith(".jpg"))
    return {n: (z[n + ".jpg"].tobytes(), z[n + ".rgb"]) for n in names}


@pytest.fixture(scope="module")
def emu(tmp_path_factory):
    """tests/jpeg_emulate.cpp: csrc/jpeg_core.h + csrc/jpeg_parse.cpp compiled by g++ (no CUDA), one loop per kernel."""
    out = str(tmp_path_factory.mktemp("emu") / "libjpeg_emu.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-Wno-unknown-pragmas", "-shared", "-fPIC", "-o", out, os.path.join(ROOT, "tests", "jpeg_emulate.cpp")],
                   check=True)
    lib = C.CDLL(out)

    def decode(data, shape, threads=0):
        rgb = np.zeros(shape, np.uint8)
        nseg = C.c_int32()
        rc = lib.emu_jpeg_decode(data, C.c_int64(len(data)), rgb.ctypes.data_as(C.c_void_p), C.byref(nseg), threads)
        return rc, rgb, nseg.value
    decode.rounds = lib.emu_last_rounds
    return decode


def test_oracle_matches_libjpeg_fixtures(oracle, cases):
    """Pins oracle/jpeg_oracle.c: every fixture decodes to the bytes libjpeg-turbo (through Pillow) produced."""
    checked = 0
    for name, (data, want) in cases.items():
        if name.startswith("refuse_"):
            with pytest.raises(ValueError) as e:
                oracle.jpeg_decode(data)
            assert e.value.args[0] == 12
            continue
        h, w, c = oracle.jpeg_info(data)
        assert (h, w) == want.shape[:2] and c == (1 if name.startswith("gray") else 3)
        assert np.array_equal(oracle.jpeg_decode(data), want), name
        checked += 1
    assert checked >= 50


def test_kernel_functions_on_cpu_match_fixtures(emu, cases):
    """The code the kernels run per thread (Huffman chunk, DC run, 8x8 block, pixel) and the chunk-parallel decoding
    scheme around it, bit-exact on every fixture; files with restart markers split into several segments.  CTA sizes
    of 1 and 2 chunks force the multi-wave path and the longest correction chains."""
    for threads in (0, 2, 1):
        for name, (data, want) in cases.items():
            rc, rgb, nseg = emu(data, want.shape, threads)
            if name.startswith("refuse_"):
                assert rc == 12
                continue
            assert rc == 0, name
            assert np.array_equal(rgb, want), (name, threads)
            if "rst" in name:
                assert nseg > 1, name
            else:
                assert nseg == 1, name


def test_kernel_functions_on_cpu_match_oracle_on_photo_sized_image(emu, oracle):
    from PIL import Image
    import io
    rng = np.random.default_rng(5)
    y, x = np.mgrid[0:480, 0:640]
    img = np.clip(np.stack([128 + 100 * np.sin(x / 37.0 + c) * np.cos(y / 23.0) for c in range(3)], -1)
                  + rng.normal(0, 12, (480, 640, 3)), 0, 255).astype(np.uint8)
    b = io.BytesIO()
    Image.fromarray(img).save(b, "JPEG", quality=85)
    data = b.getvalue()
    want = oracle.jpeg_decode(data)
    for threads in (0, 64):                        # one wave of 1024 chunks; waves of 64
        rc, rgb, _ = emu(data, (480, 640, 3), threads)
        assert rc == 0
        assert np.array_equal(rgb, want)
    # a wrongly started decoder falls into step quickly: far fewer correction rounds than chunks
    emu(data, (480, 640, 3), 0)
    assert emu.rounds() < 40 < len(data) * 8 // 1024


def test_jpeg_info_and_header_errors(ftf, cases):
    from flink_tensorflow_b200 import device
    data, want = cases["420_17x33_q75"]
    assert device.jpeg_info(data) == (17, 33, 3)
    assert device.jpeg_info(cases["gray_9x70"][0]) == (9, 70, 1)
    with pytest.raises(NotImplementedError):                       # UNIMPLEMENTED -> UnsupportedOperationException
        device.jpeg_info(cases["refuse_progressive"][0])
    for bad in (b"", b"\xff", b"not a jpeg at all", data[:20], data[:2] + b"\xff\xd9"):
        with pytest.raises(ValueError):                            # INVALID_ARGUMENT -> IllegalArgumentException
            device.jpeg_info(bad)
    # a table the scan needs is missing: drop the first DHT segment
    i = data.index(b"\xff\xc4")
    ln = int.from_bytes(data[i + 2:i + 4], "big")
    with pytest.raises(ValueError):
        device.jpeg_info(data[:i] + data[i + 2 + ln:])


def test_corrupt_entropy_data_is_reported_by_the_emulated_kernel(emu, cases):
    data, want = cases["444_opt"]
    # an optimised table has unused long codes: a run of 1 bits is not a code
    i = data.index(b"\xff\xda")
    ln = int.from_bytes(data[i + 2:i + 4], "big")
    start = i + 2 + ln
    broken = data[:start] + b"\xfe" * 64 + data[start + 64:]
    rc, _, _ = emu(broken, want.shape)
    assert rc in (0, 3)          # either decodes as garbage or trips the invalid-code check; must not crash


def test_damaged_files_never_hang_or_crash_the_kernel_functions(emu, cases, ftf):
    """Random damage to headers and entropy-coded bytes: the host reader refuses or accepts, and whatever it accepts the
    (emulated) kernels decode to completion -- every loop of decode_chunk advances, the correction rounds are bounded by
    the number of chunks."""
    from flink_tensorflow_b200 import device
    rng = np.random.default_rng(99)
    picks = ["420_opt", "444_rst_blocks", "422_q30", "gray_rst", "420_40x41_q75"]
    decoded = refused = 0
    for name in picks:
        data, want = cases[name]
        sos = data.index(b"\xff\xda")
        for trial in range(40):
            b = bytearray(data)
            lo = sos if trial % 2 else 2                       # odd trials: entropy data only; even: anywhere
            for _ in range(int(rng.integers(1, 6))):
                b[int(rng.integers(lo, len(b)))] = int(rng.integers(0, 256))
            if trial % 7 == 0:
                b = b[: int(rng.integers(lo + 1, len(b)))]     # and truncated
            b = bytes(b)
            try:
                h, w, _ = device.jpeg_info(b)
            except (ValueError, NotImplementedError):
                refused += 1
                continue
            if h * w > 1 << 20:                                # a damaged size field: keep the test small
                continue
            rc, rgb, _ = emu(b, (h, w, 3))
            assert rc in (0, 3)
            decoded += 1
    assert decoded > 100 and refused > 0


def test_decoder_has_no_cpu_path(ftf):
    """Without a CUDA device the batch decoder cannot even be created: FAILED_PRECONDITION, loudly."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from flink_tensorflow_b200 import _abi
    lib = _abi.load_library()
    h = C.c_void_p()
    rc = lib.ftf_jpeg_batch_create(0, C.byref(h))
    assert rc == _abi.FAILED_PRECONDITION and not h.value
    assert b"no CPU path" in lib.ftf_jpeg_batch_error(None)
    lib.ftf_jpeg_batch_destroy(None)                          # NULL-safe
