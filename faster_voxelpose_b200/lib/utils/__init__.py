"""Host-side helpers mirrored from reference lib/utils where the modules need them."""
