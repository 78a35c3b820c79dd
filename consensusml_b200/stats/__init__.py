"""stats"""
