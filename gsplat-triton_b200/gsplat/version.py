__version__ = "1.4.0+b200"
