"""The byte / cycle model behind DESIGN.md section 4.3: for the three 64->64 kernels at a given frame size, the
tensor-pipe floor, the shared-memory operand traffic of the instruction stream, the HBM bytes, and the bound each
implies; then the same for the variants that were considered (CTA pairs, tap stacking in N, TMA-fed weight
gradient).  Pure arithmetic, no GPU:   python tools/model_conv64.py [H W] [clock_GHz]

Constants: 148 SMs; shared memory 128 B/clk/SM; tcgen05 kind::f16 floor max(M,128) * N / 256 cycles per K = 16
instruction; measured HBM copy bandwidth and sustained tensor peak from MEASURED_PEAKS.json (fallbacks as in
frame2frame_b200/benchutil.py)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from frame2frame_b200.benchutil import peaks

H, W = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1080, 1920)
GHZ = float(sys.argv[3]) if len(sys.argv) > 3 else 1.7
SMS, SMEM_BPC = 148, 128
pk, kind = peaks()
P = H * W
ACT = P * 128                                   # one NHWC 16-bit activation tensor, bytes
FLOP = 2 * 9 * 64 * 64 * P


def mma_cycles(M, N):
    return max(M, 128) * N / 256.0


def us(cycles_per_sm):
    return cycles_per_sm / (GHZ * 1e3)


def report(name, tiles, mma_list, other_smem_per_tile, hbm_bytes, measured_us=None):
    """mma_list: [(count per tile, M, N)]; operand bytes per instruction = (M + N) * 16 * 2."""
    per_sm = tiles / SMS
    tensor = sum(m[0] * (m[3] if len(m) > 3 else mma_cycles(m[1], m[2])) for m in mma_list)    # optional explicit cycles
    smem = sum(m[0] * (m[1] + m[2]) * 32 for m in mma_list) + other_smem_per_tile
    t_tensor, t_smem, t_hbm = us(per_sm * tensor), us(per_sm * smem / SMEM_BPC), hbm_bytes / pk["hbm_gbs"] / 1e3
    print("%-44s tensor %6.1f us | smem %4.0f KB/tile -> %6.1f us | HBM %5.0f MB -> %6.1f us | bound %6.1f us%s"
          % (name, t_tensor, smem / 1024, t_smem, hbm_bytes / 1e6, t_hbm, max(t_tensor, t_smem, t_hbm),
             "" if measured_us is None else " | measured %.0f us" % measured_us))


print("frame %dx%d, %.2f GHz, HBM %.0f GB/s (%s), %.1f GFLOP per launch = %.1f us at the sustained tensor peak"
      % (W, H, GHZ, pk["hbm_gbs"], kind, FLOP / 1e9, FLOP / pk.get("bf16_tflops_sustained", pk["bf16_tflops"]) / 1e6))
t128 = -(-W // 8) * -(-H // 16)                 # k_conv64t tiles: 8 x 16 pixels
halo = 180 * 128
print("\n-- as built")
# forward: TMA tile 23 KB in, in-place BN+ReLU transform (read + write), epilogue staging 2 x 16 KB, 36 MMAs M128 N64
report("forward (BN+ReLU prologue, statistics)", t128, [(36, 128, 64)], halo * 3 + 2 * 16384, 2 * ACT, 156)
# input gradient: raw TMA tile, pre-activation tile for the epilogue, staging
report("input gradient (+ BatchNorm-backward sums)", t128, [(36, 128, 64)], halo + 16384 + 2 * 16384, 3 * ACT, 145)
# weight gradient: tiles of 8 rows x 16 px; per output row 3 MMAs M128 N64 + 3 MMAs M64 N64; producers store x and dz tiles
twg = -(-W // 16) * -(-H // 8)
x_halo = 10 * 18 * 128
report("weight gradient (register-staged operands)", twg, [(24, 128, 64), (24, 64, 64)], x_halo + 16384,
       ACT * (1.406 + 3), 218)
print("\n-- variants")
# CTA pair: each SM reads its A and half of B locally; the other half arrives from the neighbour (not through this port)
report("forward as a CTA pair (per SM, local reads only)", t128, [(36, 128, 32, 32.0)], halo * 3 + 2 * 16384, 2 * ACT, 207)
print("   measured: 48 cycles per M256 instruction per pair = per SM no faster than 44 for M128 N64 alone")
# weight gradient with the three dx taps stacked in N (B = three / four halo rows at the layout's chunk stride)
cols_pair = 3 * 192                              # accumulators: t = 1, 2 full, t = 0, 3 half-used (M = 128 pairs of dz rows)
print("weight gradient, taps stacked in N = 192: operand reads 18 KB instead of 30 KB per 16 pixels, but %d TMEM columns "
      "(> 512); with M = 64 instructions 3 x 192 = 576 columns at half the tensor rate" % cols_pair)
# TMA-fed weight gradient: TMA writes x, g, z tiles; in-place transforms; same MMAs
tma_other = (x_halo + 2 * 16384) + (2 * x_halo + 2 * 16384 + 16384)
report("weight gradient, TMA-fed + in-place transforms", twg, [(24, 128, 64), (24, 64, 64)], tma_other, ACT * (1.406 + 3))
report("  ... with post-ReLU a_k stored by the forward", twg, [(24, 128, 64), (24, 64, 64)],
       (x_halo + 2 * 16384) + (2 * 16384 + 16384), ACT * (1.406 + 3))
print("\nper training iteration (14 + 14 + 15 launches): HBM %.1f GB -> %.2f ms; tensor floor %.2f ms"
      % ((14 * 2 + 14 * 3 + 15 * 4.406) * ACT / 1e9, (14 * 2 + 14 * 3 + 15 * 4.406) * ACT / pk["hbm_gbs"] / 1e6,
         43 * FLOP / pk.get("bf16_tflops_sustained", pk["bf16_tflops"]) / 1e9))
