#!/bin/bash
# Site settings sourced by the launchers (the reference ships an empty settings.sh too).
# export B200PS_HDFS_ROOT=/mnt/hdfs_mirror   # where hdfs:// URIs resolve locally
