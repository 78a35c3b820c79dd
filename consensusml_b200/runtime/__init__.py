"""Runtime: configuration tree, artefact store / resume, run log, timers, failure checks, clock sampling."""
