"""Re-export of the seeded synthetic-input generators (they are shared with the product's
bench so both arms see identical inputs)."""
from sam2_opt_b200.synth import *  # noqa: F401,F403
from sam2_opt_b200.synth import checksum  # noqa: F401
