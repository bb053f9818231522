"""Logging, timing, clock sampling."""
