"""librectify_b200 — B200-native hot path of librectify behind the reference's C API."""
