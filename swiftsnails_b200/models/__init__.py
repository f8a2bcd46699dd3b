"""swiftsnails_b200.models"""
