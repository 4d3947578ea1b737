"""Command-line utilities (python -m distributed_tensorflow_framework_b200.cli.<name>)."""
