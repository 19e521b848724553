"""Modules with the reference lib/models interfaces, backed by libfvp_b200."""
