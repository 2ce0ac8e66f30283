"""Mirrors of the reference's util/ helpers that sit on the update path."""
