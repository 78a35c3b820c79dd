"""Consensus layer: set intersections, rank aggregation, the per-gene standard table, gene-list writers."""
from .gene_csv import write_gene_list, write_lasso_gene_csv  # noqa: F401
from .rank_aggregate import aggregate, top_k_overlap  # noqa: F401
from .sets import intersections, membership_table, selection_sets  # noqa: F401
from .standtable import StandTable  # noqa: F401
