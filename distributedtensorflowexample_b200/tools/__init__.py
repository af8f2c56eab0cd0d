"""Command-line tools a TF1 user expects next to the trainer: ``inspect_checkpoint`` and ``read_events``."""
