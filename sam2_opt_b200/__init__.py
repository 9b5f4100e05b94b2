"""sam2_opt_b200 — B200-native (sm_100a) SAM2.1 Hiera-L inference hot path behind the
reference's ``predictor.speedup()`` hook.  See DESIGN.md."""
__version__ = "0.1.0"
