"""Empty stand-in for `torchgeometry`: the reference imports it (utils/loss_utils.py:17, `import torchgeometry as
tgm`) and never touches the name again on the training path, so an importable module is all it needs."""
