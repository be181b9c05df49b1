#!/usr/bin/env bash
# Offline self-check (no network, no GPU needed): imports, tier A/B models build on the meta device with
# the documented parameter counts, the synthetic dataset has the right shape, config files parse, the
# native library builds for sm_100a.  Mirrors what the reference's verify_offline.sh checks inside
# `docker run --network none`; pass --docker IMAGE to run the same checks inside an image that way.
set -euo pipefail
cd "$(dirname "${BASH_SOURCE[0]}")/.."
if [ "${1:-}" = "--docker" ]; then
  exec docker run --rm --network none --entrypoint bash "${2:?image}" -c "cd /app && scripts/verify_offline.sh"
fi
python - <<'PY'
import importlib, sys
sys.path.insert(0, ".")
print("[1/5] imports")
for m in ("torch", "pandas", "yaml", "PIL", "b200bench", "b200bench.cli", "b200bench.engine", "b200bench.analysis.plot"):
    importlib.import_module(m); print("      ok", m)
import torch
from b200bench.config import get_model_config, load_deepspeed_config, load_fsdp_config
from b200bench.models import TinyGPT, build_layout
print("[2/5] models (meta device)")
for tier, seq, want in (("A", 2048, 236_406_784), ("B", 2048, 1_681_199_104)):
    with torch.device("meta"):
        m = TinyGPT(verbose=False, **get_model_config(tier, seq))
    n = m.num_params(); assert n == want == build_layout(**get_model_config(tier, seq)).num_params(), (tier, n)
    print(f"      ok tier {tier}: {n/1e6:.2f}M parameters")
print("[3/5] synthetic dataset")
from b200bench.utils.data import SyntheticDataset
ds = SyntheticDataset(32000, 2048, size=100, verbose=False)
assert tuple(ds.data.shape) == (100, 2048) and int(ds.data.max()) < 32000; print("      ok", tuple(ds.data.shape))
print("[4/5] config files")
for s in ("zero2", "zero3"):
    c = load_deepspeed_config(f"configs/deepspeed/{s}.json", 1, 4, 8); assert c.raw["train_batch_size"] == 32
    print("      ok", s, "clip", c.optim.grad_clip, "warmup", c.optim.warmup_steps)
print("      ok fsdp", load_fsdp_config("configs/fsdp/fsdp_config.yaml").fsdp_sharding_strategy)
PY
echo "[5/5] native build (nvcc, sm_100a)"
if command -v nvcc > /dev/null; then make -j8 > /dev/null && echo "      ok $(ls b200bench/lib/libb200bench.so)"; else echo "      skipped: no nvcc in this environment"; fi
echo "offline verification passed"
