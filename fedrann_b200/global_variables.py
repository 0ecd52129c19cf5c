"""Run-wide state, same names as fedrann/global_variables.py:4-7 (threads, seed,
output_dir, temp_dir).  `threads` only sizes host-side file parsing here."""
threads: int = 1
seed: int = 356115
output_dir: str = ""
temp_dir: str = ""
