"""Empty matplotlib stand-in: the reference imports it from out-of-scope envs only."""
