"""I/O layer: R object reader, reference-format CSV readers/writers, count-file ingestion."""
