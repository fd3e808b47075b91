"""CPU oracle — TEST INFRASTRUCTURE ONLY.  See oracle/orc_common.h for the provenance header ("parity unpinned")."""
