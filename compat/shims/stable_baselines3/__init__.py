"""Import stand-in for stable-baselines3 1.7.0.  The reference imports these names (reference
src/rlsp/agents/simple_ddpg.py:17, src/rlsp/agents/main.py:21-22) but uses them only when
``graph_mode`` is off, which the GNN agent never is."""
