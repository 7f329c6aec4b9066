"""Reference-arm harness: runs the UNMODIFIED reference (baseline/_ref or /root/reference)."""
