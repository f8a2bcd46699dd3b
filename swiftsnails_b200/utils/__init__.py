"""swiftsnails_b200.utils"""
