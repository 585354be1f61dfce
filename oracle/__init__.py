"""Parity oracle: TEST INFRASTRUCTURE.  See bdilu_oracle.c / ref_shim.cpp / loader.py headers."""
