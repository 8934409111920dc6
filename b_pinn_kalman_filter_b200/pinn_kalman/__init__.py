"""Callers of the Navier-Stokes step and of grid_sample on the B200 ops (reference pinn_kalman/, models/flownet.py)."""
