"""swiftsnails_b200.parallel"""
