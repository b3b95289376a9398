"""Import-compatibility package: the reference is imported as ``src.*``
(``packages = [{include = "src"}]`` in its pyproject).  Everything here
re-exports the implementation in ``distributed_training_b200``."""
