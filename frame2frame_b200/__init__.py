"""frame2frame_b200 -- the frame2frame hot path on B200 (sm_100a): host-side mirrors of the reference's Python
interface over libf2f_b200.so (include/f2f_b200.h).

    blind_denoising   the reference's CLI / `blind_denoising(**args)` / `psnr` + the OnlineF2F engine
    warped_loss       WarpedLoss (paper script) and LibWarpedLoss (lib/frame2frame flavour)
    dncnn             DnCNN, Adam
    tvl1, tvl1flow_cli    tvl1flow() and the two TV-L1 command lines, .flo I/O
    batch_f2f         batch-mode fine-tuning with the NCCL gradient exchange
    lib_api           frame2frame.load_model / extract_config (also importable as `import frame2frame`)

There is no CPU fallback: importing is free, every compute call needs the built library and a CUDA device.
"""
