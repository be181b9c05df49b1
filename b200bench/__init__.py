"""b200bench — Blackwell-native distributed-training benchmark framework.

Same capabilities as deepaksatna/Distributed-LLM-Training-Benchmark-Framework (one
``train_harness``-style CLI, ``--strategy {ddp,fsdp,zero2,zero3}`` on the TinyGPT decoder, the
13-key result JSON, CSV / plots / markdown report), rebuilt for one 8xB200 NVSwitch box:
hand-written sm_100a kernels (``csrc/``) driven by a native engine, with PyTorch as the
framework layer and ``torch.distributed`` for rendezvous.

Layout
    models/    TinyGPT (torch oracle with the reference's state_dict), flat parameter layout
    ops/       loader + Python front-ends of the native kernels (``libb200bench.so``)
    engine.py  native training engine: manual forward/backward over flat buffers
    parallel/  DDP / FSDP / ZeRO-2 / ZeRO-3 strategy objects, peer-memory runtime
    optim/     flat AdamW (fp32 master + bf16 working weights), WarmupLR, clipping
    utils/     timers, result schema, clocks sampling, synthetic data
    analysis/  parse_metrics / plot / make_report (reference CLIs and file names)
"""

__version__ = "0.1.0"
