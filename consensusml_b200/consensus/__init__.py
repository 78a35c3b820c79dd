"""consensus"""
