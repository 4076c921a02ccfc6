"""Import-path compatibility with the reference: ``pytorch_robotics_transformer.*`` resolves to ``rt1b200.models``."""
