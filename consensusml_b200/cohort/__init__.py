"""cohort"""
