"""Worker-side pieces of the decode path (reference: worker/)."""
