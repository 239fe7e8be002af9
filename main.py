"""`python main.py ...` -- same command line as the reference's main.py; the work happens in
distributed_pytorch_b200.cli.run("single")."""
from distributed_pytorch_b200.cli import main

if __name__ == "__main__":
    main("single")
