"""Native parallel engines: symmetric memory, data parallel, sharded data parallel, TP / SP / CP."""
