"""CPU oracle (test infrastructure). Importable only from tests/, __graft_entry__.smoke()
and bench.py's CPU-baseline legs — never from the product package."""
