"""B200-native GPT training framework with the capabilities of KT19/distributed-training-compare-jax.

Layers (top to bottom):
  main.py / plot.py / bench.py      CLI entry points (same ``--train_config_path`` contract as the reference)
  train/                            step builders, optimizer factory, training drivers (DP / TP / PP)
  parallel/                         device-mesh runtime, sharding rules, DP / TP / PP engines, symmetric memory
  models/                           GPTModel / TransformerBlock / CausalSelfAttention / MLP + Flax-compatible param tree
  ops/                              native sm_100a kernels (ctypes binding of libdtc_b200.so) + fp32 torch oracle
  data/, config/, utils/            synthetic FineWeb-edu-shaped stream, YAML schema, logging / timing / checkpoint
  csrc/                             hand-written CUDA (tcgen05 / TMEM / TMA) sources of libdtc_b200.so
"""
__version__ = "0.1.0"
