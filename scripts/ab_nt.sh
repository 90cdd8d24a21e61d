for r in 1 2; do
  echo "default run $r"; python scripts/dsweep_profile.py c1 device 5 | grep -v '"rep": [01],' | cut -c1-90
  echo "cluster8 run $r"; TCI_DSWEEP_CLUSTER=8 python scripts/dsweep_profile.py c1 device 5 | grep -v '"rep": [01],' | cut -c1-90
done
