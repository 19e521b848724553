"""Drop-in mirror of the reference tree layout (lib/models, lib/core, lib/utils) for the hot-path modules."""
