"""rnnt-transducer_b200: B200-native RNN-T joint network + transducer loss.

Import it as ``rnnt_transducer_b200`` (see the alias package at the repo root)."""
