"""Drop-in ``xtomosim`` package (the reference imports it at /root/reference/main.py:19-21, plots.py:16-17)."""
