"""Stand-in for ``external/getcontacts/contact_calc`` (same module names for the pieces ChACRA's path imports)."""
