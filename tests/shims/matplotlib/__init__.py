"""Headless no-op stand-in for matplotlib, used ONLY by tests that run the reference's main.py unchanged
(/root/reference/main.py:27,31-53 imports pyplot and forces text.usetex; neither exists on these machines)."""
rcParams = {}
