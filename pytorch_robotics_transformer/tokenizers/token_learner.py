from rt1b200.models.token_learner import TokenLearnerModule  # noqa: F401
