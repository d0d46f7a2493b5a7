"""core — B200-native replacements for the modules of the reference's ``core/`` package
that sit on the paged-KV attention hot path (same import paths, same public names)."""
