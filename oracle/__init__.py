"""ORACLE package: CPU restatements of the reference algorithms.  Test infrastructure only --
see the header of each module.  Never imported by yolo_series_b200/."""
