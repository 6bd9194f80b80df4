"""CPU oracle for the SHA-256 content-addressing path.

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and
bench.py's CPU-baseline legs, never from bee_code_interpreter_fs_b200.
"""
