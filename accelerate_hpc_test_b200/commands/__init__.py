"""Command line interface: ``accelerate config | launch | env | test | estimate-memory | merge-weights | to-fsdp2``."""
