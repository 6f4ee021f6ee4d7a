"""CPU oracle (test infrastructure).  Importable only from tests/, bench.py's cpu_baseline /
--impl reference legs and __graft_entry__.smoke()."""
