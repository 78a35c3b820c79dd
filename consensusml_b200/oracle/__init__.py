"""oracle"""
