"""Empty stand-in: the reference imports pygame at module top
(minigrid/minigrid_env.py:10-11) but only uses it for render_mode="human"."""
