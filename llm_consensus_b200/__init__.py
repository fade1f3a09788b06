"""Importable alias of the product package directory `llm-consensus_b200/` (a hyphen cannot
appear in a Python module name).  `import llm_consensus_b200.llmc` resolves to
`llm-consensus_b200/llmc.py`."""
import os as _os

__path__.insert(0, _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "llm-consensus_b200"))
