"""Algorithmic byte model of the hot path — implementation-independent bytes each kernel's function
must move, 4-byte elements. N = pixels per frame, K = keylines per frame (measured mean), B = lanes per launch,
P = measured mean directed-match probes per keyline.

Per frame:  B_frame = 64 N + (1260 + 4 P) K   (scale space 48 N, detect 12 N, distance-field reset 4 N, K terms below)
Per launch of each kernel (what `roofline.achieved` in bench.py divides by that kernel's CUDA-event time):"""
from __future__ import annotations


def frame_bytes(N, K, P=20.0):
    return 64.0 * N + (1260.0 + 4.0 * P) * K


def kernel_bytes(name, N, K, B, planes=1, P=20.0, words=0):
    """algorithmic bytes of ONE launch of kernel `name` over B lanes (planes = image planes per lane in that launch)"""
    per_lane = {
        # one box pass (createIntegralImage + average) must read N and write N floats per plane; it is split in two
        # kernels here, each of which reads N and writes N of its own
        "k_rowscan": 8.0 * N * planes,
        "k_rowscan4": 8.0 * N * planes,
        "k_colbox": 8.0 * N * planes,
        "k_rowscan_lat": 8.0 * N * planes,
        "k_colbox_lat": 8.0 * N * planes,
        "k_detect_flags": 12.0 * N,          # read both scales, write the id mask (DoG / magnitude never reach memory)
        "k_detect_scan": 0.0,
        "k_detect_emit": 84.0 * K,           # one KeyLine per keyline
        "k_detect_emit_wide": 84.0 * K,
        "k_join": 20.0 * K,
        "k_tune": 4.0 * K,
        "k_field_build": 340.0 * K,          # 20 B keyline read + 80 candidates x 4 B
        "k_field_band": 4.0 * N + 340.0 * K,  # same + the field reset/write it absorbs (4 N)
        "k_boxpass": 8.0 * N * planes,       # one whole box pass: read N + write N per plane
        "k_boxpass2": 8.0 * N * planes,
        "k_pack_targets": 52.0 * K,          # 20 B read + 32 B packed write
        "k_rotate": 48.0 * K,
        "k_quantile": 4.0 * K,
        "k_tryvel_fast": 68.0 * K, "k_tryvel_fast2": 68.0 * K, "k_tryvel_fast4": 68.0 * K,
        "k_tryvel_ord2": 68.0 * K, "k_tryvel_ord4": 68.0 * K, "k_minimize_ord2": 6 * 68.0 * K, "k_minimize_ord4": 6 * 68.0 * K,   # the addend planes the chain kernel reads back (40 B written + 40 B read per keyline) are not algorithmic
        "k_chain_tryvel": 0.0, "k_chain_extrot": 0.0,           # serial float32 add chains: latency, not bytes
        "k_extrot_fast": 40.0 * K, "k_extrot_ord": 40.0 * K,
        "k_tryvel_eval": 68.0 * K,           # 36 B keyline + 4 B field + 20 B target + 8 B write
        "k_tryvel_reduce": 48.0 * K,         # 10 addends + residual code read, residual write
        "k_fwd_link": 12.0 * K,
        "k_fwd_resolve": 60.0 * K,
        "k_extrot_rows": 40.0 * K,
        "k_extrot_reduce": 28.0 * K,
        "k_directed": (72.0 + 4.0 * P) * K,
        "k_reg_eval": 60.0 * K,
        "k_reg_apply": 8.0 * K,
        "k_kalman": 52.0 * K,
    }
    if name == "k_fill":
        return 4.0 * words
    return per_lane.get(name, 0.0) * B
