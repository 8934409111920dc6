"""Inverse-problem samplers on the B200 score engine (reference inverse/): DPS with the network's input gradient."""
