"""The legs of bench.py, one module per responsibility:
  clocks      nvidia-smi clock/throttle sampling during the timed region
  data        synthetic sample batches (seeded; numpy for cfg2 so that every arm sees the same stream)
  cpu_ref     the reference's CPU code timed on the host cores, the reference arm, the parity block
              (the ONLY bench module that touches oracle/)
  em_leg      the headline: device-resident training steps, roofline of the E-step kernel
  e2e_leg     the same step through host buffers (sdmmb_optimize_host)
  query_leg   guided sample+pdf queries (second half of BASELINE.json's metric)
  cfg5_leg    BASELINE.json configs[4]: 65536 x 128 x 8192 sharded by cell, timed in every --gpus N run
  render_leg  progressive training loop (configs[0] / [3])
"""
