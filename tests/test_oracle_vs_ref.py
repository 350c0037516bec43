"""Pins the oracle (oracle/paged_attn_oracle.c) against the reference's own ggml CPU code (oracle/_ref).

The reference's tests hold no golden vectors for this path, so the pin is the reference
itself, compiled from /root/reference and run in this container; tests/test_golden.py re-checks the same
quantities from committed fixtures on machines without the reference tree."""
import numpy as np
import pytest

from workloads import F16, Q4_0, Q8_0, make_workload

TYPES = [F16, Q8_0, Q4_0]


def _inputs(seed, n=4096, k=128):
    rng = np.random.default_rng(seed)
    x = rng.uniform(-1, 1, (n, k)).astype(np.float32)
    x[0, :32] = 0.0                         # all-zero block: d == 0 -> id = 0
    x[1, :32] = np.float32(1e-30)           # d underflows in half
    x[2, 7] = np.float32(6.0e4)             # near half max
    x[3, :] = rng.standard_normal(k).astype(np.float32) * 100
    x[4, :32] = np.float32(-0.75)           # ties in amax, negative max for Q4_0
    x[5, :32] = np.linspace(-1, 1, 32, dtype=np.float32)
    x[5, 31] = -1.0                         # |max| tie with opposite signs: first wins
    return x


@pytest.mark.parametrize("t", TYPES)
def test_quantisers_bit_exact_with_reference_ref(oracle, ref, t):
    """quantize_row_q8_0_ref / quantize_row_q4_0_ref / ggml_fp32_to_fp16_row, byte for byte"""
    for seed in (1, 2, 3):
        x = _inputs(seed)
        assert np.array_equal(oracle.from_float(t, x), ref.quantize(t, x))


@pytest.mark.parametrize("t", TYPES)
def test_dequantisers_bit_exact(oracle, ref, t):
    x = _inputs(11)
    b = ref.quantize(t, x)
    assert np.array_equal(oracle.to_float(t, b, 128), ref.to_float(t, b, 128))


def test_half_conversion_all_bit_patterns(oracle, ref):
    """every binary16 pattern round-trips like the reference; fp32->fp16 on boundaries"""
    h = np.arange(65536, dtype=np.uint16)
    f_ref = ref.to_float(F16, h.view(np.uint8).reshape(-1, 2 * 128), 128).reshape(-1)
    f_orc = oracle.to_float(F16, h.view(np.uint8).reshape(-1, 2 * 128), 128).reshape(-1)
    assert np.array_equal(f_ref.view(np.uint32), f_orc.view(np.uint32))
    finite = f_ref[np.isfinite(f_ref)]
    # midpoints and neighbours of every finite half
    probes = np.concatenate([finite, np.nextafter(finite, np.inf), np.nextafter(finite, -np.inf),
                             (finite[:-1] + finite[1:]) / 2]).astype(np.float32)
    probes = probes[: probes.size // 128 * 128].reshape(-1, 128)
    assert np.array_equal(oracle.from_float(F16, probes), ref.quantize(F16, probes))


@pytest.mark.parametrize("t", [Q8_0, Q4_0])
def test_cpu_backend_cpy_vs_ref_quantiser(oracle, ref, t):
    """GGML_OP_CPY on the CPU backend uses its SIMD quantiser; the reference's own test_cpy only asks for
    NMSE <= 1e-6 after dequantisation (tests/test-backend-ops.cpp:1469-1471).  Count the differing codes."""
    x = _inputs(5, n=2048)
    a = ref.cpy_f32(t, x, n_threads=2)
    b = oracle.from_float(t, x)
    fa, fb = ref.to_float(t, a, 128), oracle.to_float(t, b, 128)
    assert oracle.nmse(fa, fb) <= 1e-6
    frac = float((a != b).mean())
    assert frac < 1e-4, frac


CASES = [
    # (seq_lens, n_head, n_head_kv, d)
    ([37, 1, 64, 130], 8, 2, 128),
    ([100, 17], 4, 4, 64),
    ([256], 16, 2, 128),
]


@pytest.mark.parametrize("t", TYPES)
@pytest.mark.parametrize("case", CASES)
def test_attention_oracle_vs_reference(oracle, ref, t, case):
    """paged oracle == reference CPU on the dense-mask equivalent of the same cache bytes:
    F32 formulation (mul_mat/soft_max_ext/mul_mat) to 2e-5, FLASH_ATTN_EXT op to its own NMSE bar 5e-4"""
    seq_lens, n_head, n_head_kv, d = case
    w = make_workload(oracle, seq_lens, n_head, n_head_kv, d, t, seed=1234 + t)
    got = oracle.attention(w.q, w.k_cache, w.v_cache, t, t, n_head_kv, w.scale, block_table=w.block_table,
                           seq_lens=w.seq_lens, block_size=w.block_size)
    mask = w.dense_mask()
    f32, _ = ref.attention(1, w.q, w.k_cache, w.v_cache, t, t, n_head_kv, w.n_cells, mask, w.scale, n_threads=2)
    fa, _ = ref.attention(0, w.q, w.k_cache, w.v_cache, t, t, n_head_kv, w.n_cells, mask, w.scale, n_threads=2)
    assert oracle.max_rel_err(f32, got) < 2e-5
    assert oracle.nmse(fa, got) < 5e-4
    # and the oracle's own dense-mask mode agrees with its paged mode
    dense = oracle.attention(w.q, w.k_cache, w.v_cache, t, t, n_head_kv, w.scale, mask=mask, n_kv=w.n_cells)
    assert np.array_equal(dense, got)


def test_attention_softcap_and_alibi_dense(oracle, ref):
    """dense-mask mode with logit_softcap and max_bias (ALiBi slope on the mask), test_flash_attn_ext's knobs"""
    rng = np.random.default_rng(3)
    n_q, n_head, n_head_kv, d, n_kv = 3, 8, 2, 128, 96
    q = rng.uniform(-1, 1, (n_q, n_head, d)).astype(np.float32)
    kf = rng.uniform(-1, 1, (n_kv, n_head_kv * d)).astype(np.float32)
    vf = rng.uniform(-1, 1, (n_kv, n_head_kv * d)).astype(np.float32)
    k = oracle.from_float(F16, kf.reshape(-1, d)).reshape(n_kv, -1)
    v = oracle.from_float(F16, vf.reshape(-1, d)).reshape(n_kv, -1)
    mask = rng.uniform(-1, 1, (64, n_kv)).astype(np.float16)     # finite mask like init_tensor_uniform
    mask[1, 5:20] = -np.inf
    for max_bias, softcap in ((0.0, 10.0), (8.0, 0.0), (8.0, 10.0)):
        got = oracle.attention(q, k, v, F16, F16, n_head_kv, 0.088, mask=mask, n_kv=n_kv, max_bias=max_bias, softcap=softcap)
        f32, _ = ref.attention(1, q, k, v, F16, F16, n_head_kv, n_kv, mask, 0.088, max_bias, softcap, n_threads=2)
        fa, _ = ref.attention(0, q, k, v, F16, F16, n_head_kv, n_kv, mask, 0.088, max_bias, softcap, n_threads=2)
        assert oracle.max_rel_err(f32, got) < 2e-5, (max_bias, softcap)
        assert oracle.nmse(fa, got) < 5e-4


@pytest.mark.parametrize("t", TYPES)
@pytest.mark.parametrize("hs", [64, 128])
def test_oracle_vs_reference_over_the_reference_test_grid(oracle, ref, hs, t):
    """The whole test_flash_attn_ext loop nest of tests/test-backend-ops.cpp:4367-4391 (the grid the GPU path is run over
    in tests/test_gpu_ggml_entry.py), oracle against the reference CPU backend: F32 formulation to 2e-5 (3e-5 with
    softcap: tanhf), the FLASH_ATTN_EXT op to its own NMSE bar"""
    from test_gpu_ggml_entry import _grid
    worst = [0.0, 0.0]
    for i, c in enumerate(_grid(hs)):
        if c["permuted"]:
            continue                                         # same values, other strides: the harness takes token-major caches
        rng = np.random.default_rng(4321 + i)
        nh, n_head, kv, nb = 4, 4 * c["nr"], c["kv"], c["nb"]
        q = rng.uniform(-1, 1, (nb, n_head, hs)).astype(np.float32)
        k = oracle.from_float(t, rng.uniform(-1, 1, (kv * nh, hs)).astype(np.float32)).reshape(kv, -1)
        v = oracle.from_float(t, rng.uniform(-1, 1, (kv * nh, hs)).astype(np.float32)).reshape(kv, -1)
        m = None
        if c["mask"]:
            m = rng.uniform(-1, 1, ((nb + 63) // 64 * 64, kv)).astype(np.float16)
            m[0, kv // 3: kv // 2] = -np.inf
        scale = 1.0 / np.sqrt(hs)
        got = oracle.attention(q, k, v, t, t, nh, scale, mask=m, n_kv=kv, max_bias=c["max_bias"], softcap=c["softcap"])
        f32, _ = ref.attention(1, q, k, v, t, t, nh, kv, m, scale, c["max_bias"], c["softcap"], n_threads=2)
        fa, _ = ref.attention(0, q, k, v, t, t, nh, kv, m, scale, c["max_bias"], c["softcap"], n_threads=2)
        e1, e2 = oracle.max_rel_err(f32, got), oracle.nmse(fa, got)
        worst = [max(worst[0], e1), max(worst[1], e2)]
        assert e1 < (3e-5 if c["softcap"] else 2e-5), (c, e1)
        assert e2 < 5e-4, (c, e2)
    assert worst[0] > 0.0                                    # the loop ran


@pytest.mark.parametrize("t", [F16, Q8_0, Q4_0])
def test_quantize_fns_thresholds_on_the_reference_test_data(oracle, ref, t):
    """tests/test-quantize-fns.cpp on the cache types: its synthetic vector 0.1 + 2 cos(i + offset) (:32-36, test_size
    32 * 128 :119), total quantisation error array_rmse(x, dequant(quant(x))) < 0.002 (:18, :134-141) and the quantiser
    against from_float_ref < 0.0001 (:17, :143-147) — here the second is byte equality with the reference's bytes"""
    n = 32 * 128
    x = (0.1 + 2.0 * np.cos(np.arange(n, dtype=np.float32) + np.float32(0.0))).astype(np.float32).reshape(-1, 128)
    b = oracle.from_float(t, x)
    assert np.array_equal(b, ref.quantize(t, x))
    back = oracle.to_float(t, b, 128)
    assert np.array_equal(back, ref.to_float(t, b, 128))
    rmse = float(np.sqrt(np.sum((x.astype(np.float64) - back) ** 2)) / n)          # array_rmse divides by n, not sqrt(n)
    assert rmse < 0.002, rmse
