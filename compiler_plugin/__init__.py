"""Compiler-side half of the drop-in: the match-and-replace transform spec (generated from
the kernel manifest) and the PluginSession source that installs it. See README.md here and
INTEGRATION.md at the repo root."""
