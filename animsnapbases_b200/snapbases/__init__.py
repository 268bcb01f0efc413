"""Drop-in mirror of the reference's `snapbases` package (same module, class, method and attribute names):
`from animsnapbases_b200.snapbases.posComponents import posComponents` replaces `from snapbases.posComponents import ...`."""
