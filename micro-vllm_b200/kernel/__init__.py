"""kernel — sm_100a CUDA replacements for the reference's Triton ops (same names/signatures)."""
