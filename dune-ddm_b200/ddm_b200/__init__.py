"""ddm_b200: B200-native solve phase of two-level overlapping Schwarz (dune-ddm hot path)."""
