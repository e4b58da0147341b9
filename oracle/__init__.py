"""oracle/ — TEST INFRASTRUCTURE ONLY.  CPU restatement (dindel_oracle.c) and, when built, the
reference's own code (oracle/_ref).  Importable only from tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs; the product package never imports it."""
