#!/usr/bin/env python3
"""Shared-memory-port model of the tcgen05 kernels (profiles/r01e_smem_operand_model.md), as arithmetic.

An SS-mode `tcgen05.mma.kind::f16` with M = 128 reads A (128 x 16 bf16 = 4 KB) and B (N x 16 bf16) from shared
memory for every instruction and runs N / 2 cycles; a TS-mode one reads B only.  Everything else that touches
shared memory -- TMA fills, epilogue loads / stores -- goes through the same 128 B/cycle/SM port.  Per tile:

    port cycles = bytes through the port / 128          mma cycles = sum over MMAs of N / 2

The larger of the two bounds the tile.  CPU only; prints the table for the shipped and the opt-in kernels.

    python tools/port_model.py
"""
PORT = 128.0          # B/cycle/SM (B300_MICROARCH.md: "smem crossbar BW 128/N B/cyc/SM")
KB = 1024.0


def mma(n_mmas, N, a_from_smem=True, b_from_smem=True):
    """-> (operand bytes fetched from shared memory, tensor-pipe cycles) of n_mmas M = 128, K = 16 instructions."""
    per = (4 * KB if a_from_smem else 0.0) + (N * 32.0 if b_from_smem else 0.0)
    return n_mmas * per, n_mmas * N / 2.0


def row(name, parts, other_bytes, note=""):
    fetch = sum(p[0] for p in parts)
    cycles = sum(p[1] for p in parts)
    total = fetch + other_bytes
    port = total / PORT
    bound = max(port, cycles)
    print(f"| {name} | {fetch / KB:.0f} | {other_bytes / KB:.0f} | {total / KB:.0f} | {port:.0f} | {cycles:.0f} | "
          f"{'port' if port > cycles else 'tensor'} | {100 * cycles / bound:.0f} % | {note} |")


def main():
    print("| kernel, per tile | MMA operand fetch KB | fills + epilogue KB | total KB | port cycles | MMA cycles | "
          "bound | tensor pipe at the bound | measured |")
    print("|---|---|---|---|---|---|---|---|---|")
    # objective, KP = 128, one V plane, 128 x 128 tile: 2 k-chunks x 4 k-steps x 3 products; fills: H^T 64 KB + V 32 KB;
    # epilogue reads V 32 KB
    row("residual_kernel<2,1> 128x128", [mma(24, 128)], (64 + 32 + 32) * KB, "2870 cycles (r01c, 1605 MHz)")
    row("residual_n256_kernel<2> 128x256 (opt-in)", [mma(24, 256)], (128 + 64 + 64) * KB, "per 2 tiles of the above")
    row("residual_ts_kernel<2,1> 128x128 (opt-in)", [mma(24, 128, a_from_smem=False)], (64 + 32 + 32) * KB, "")
    # two V planes (ADMM / AO-ADMM objective after a fixed-point ingest): V fills and reads double
    row("residual_kernel<2,2> 128x128", [mma(24, 128)], (64 + 64 + 64) * KB, "")
    row("residual_ts_kernel<2,2> 128x128 (opt-in)", [mma(24, 128, a_from_smem=False)], (64 + 64 + 64) * KB, "")
    # fused KL pass, KP = 128, 128 x 64 tile: MMA1 24 x (N = 64), MMA2 12 x (N = 128); Q stores 32 KB, Y fill 32 KB,
    # D fill 16 KB, D reads 16 KB
    row("kl_panel_kernel<2,1> 128x64", [mma(24, 64), mma(12, 128)], (32 + 32 + 16 + 16) * KB,
        "<= 3300 cycles (r01c H side, 1965 MHz)")
    row("kl_panel_ts_kernel<2,1> (opt-in)", [mma(24, 64, a_from_smem=False), mma(12, 128)], (32 + 32 + 16 + 16) * KB, "")
    row("kl_panel_ts2_kernel<2,1> (opt-in)", [mma(24, 64, a_from_smem=False), mma(12, 128, a_from_smem=False)],
        (32 + 16 + 16) * KB, "")
    # exact-hi GEMM, per 64-deep block of a 128 x 128 tile: 4 k-steps x 5 products; fills 3 factor + 2 data planes
    row("panel_gemm_exact_kernel<128,2,128> per block", [mma(20, 128)], 80 * KB, "tensor pipe 67.9 % (ncu r01b)")
    # split-bf16 panel GEMM, per 64-deep block of a 128 x 256 tile: 4 k-steps x 2 products; fills 2 factor planes
    # (multicast across the cluster: each CTA's port still receives both) + 1 data plane of 256 rows
    row("panel_gemm_kernel<128,2,1,256> per block", [mma(8, 256)], (32 + 32) * KB, "tensor pipe 85.2 % (ncu r01b)")


if __name__ == "__main__":
    main()
