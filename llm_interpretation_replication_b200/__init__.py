"""B200-native relative-probability scorer + bootstrap (hot path of
jonathanhchoi/llm-interpretation-replication).  See DESIGN.md."""
__version__ = "0.1.0"
