"""TEST INFRASTRUCTURE ONLY -- the deterministic synthetic checkpoint and synthetic views used by the
oracle, the golden-vector generator and the parity tests.  The generator itself carries no
reference arithmetic (it only needs the reference's state-dict key space), so it lives with the
product in mapanything_b200/synthetic.py, where bench.py can use it without touching oracle/."""
from mapanything_b200.synthetic import (  # noqa: F401
    add_multimodal_inputs,
    canonical_key,
    load_manifest,
    synthetic_state_dict,
    synthetic_views,
)
