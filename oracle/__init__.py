"""Test infrastructure only (CPU oracle + reference harness). Never imported by k3m_b200."""
