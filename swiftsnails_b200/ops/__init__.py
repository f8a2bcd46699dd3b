"""swiftsnails_b200.ops"""
